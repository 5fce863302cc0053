// Minimal stand-in for <sigc++/sigc++.h> (sigc++-2.0 is absent here): signal<void(Args...)> with connect() / emit()
// and ptr_fun(), as r820_dev.hpp:139 and sdrx.cpp:1772 use them.
#pragma once
#include <functional>
#include <utility>
#include <vector>
namespace sigc {
template <class Sig> class signal;
template <class R, class... A>
class signal<R(A...)> {
public:
    void connect(std::function<R(A...)> f) { slots_.push_back(std::move(f)); }
    void emit(A... a) const { for (auto &s : slots_) s(a...); }
    void operator()(A... a) const { emit(a...); }
    bool empty() const { return slots_.empty(); }
private:
    std::vector<std::function<R(A...)>> slots_;
};
template <class R, class... A> std::function<R(A...)> ptr_fun(R (*f)(A...)) { return f; }
}  // namespace sigc
