// sdrx_b200 -- specialised channel bank (placeholder until the sm_100a kernel lands).
#pragma once
#include "common.cuh"

namespace sdrx {

struct FastBank {
    bool enabled = false;
    int  stages_fused = 0;     // how many leading MSD stages the kernel covers
    void destroy() {}
};

}  // namespace sdrx
