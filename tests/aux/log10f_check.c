/* TEST AID: the C twin of sdrx_b200/csrc/common.cuh:log10f_glibc, compared with the host libm's log10f
 * (what the reference's squelch calls, sdrx.cpp:566).  argv[1] = stride over the positive floats
 * (1 = exhaustive, ~40 s).  Prints the number of mismatches. */
#include <stdlib.h>
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
static const double T[16][2] = {
  { 0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2 },
  { 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2 },
  { 0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2 },
  { 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3 },
  { 0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3 },
  { 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3 },
  { 0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4 },
  { 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4 },
  { 0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5 },
  { 0x1p+0, 0x0p+0 },
  { 0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5 },
  { 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4 },
  { 0x1.b2036576afce6p-1, 0x1.526e57720db08p-3 },
  { 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3 },
  { 0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2 },
  { 0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2 },
};
static const double Ln2 = 0x1.62e42fefa39efp-1;
static const double A[3] = { -0x1.00ea348b88334p-2, 0x1.5575b0be00b6ap-2, -0x1.ffffef20a4123p-2 };
static inline uint32_t asuint(float f){uint32_t u; memcpy(&u,&f,4); return u;}
static inline float asfloat(uint32_t u){float f; memcpy(&f,&u,4); return f;}
#ifdef USE_FMA
#define MA(a,b,c) fma((a),(b),(c))
#else
#define MA(a,b,c) ((a)*(b)+(c))
#endif
float my_logf(float x){
  uint32_t ix = asuint(x); if (ix == 0x3f800000) return 0;
  uint32_t tmp = ix - 0x3f330000; int i = (tmp >> 19) % 16; int k = (int32_t)tmp >> 23;
  uint32_t iz = ix - (tmp & 0x1ff << 23);
  double invc = T[i][0], logc = T[i][1];
  double z = (double)asfloat(iz);
  double r = MA(z, invc, -1.0);
  double y0 = MA((double)k, Ln2, logc);
  double r2 = r*r;
  double y = MA(A[1], r, A[2]);
  y = MA(A[0], r2, y);
  y = MA(y, r2, (y0 + r));
  return (float)y;
}
float my_log10f(float x){
  int32_t hx = (int32_t)asuint(x); int32_t k = 0;
  if (hx < 0x00800000) { k -= 25; x *= 3.3554432000e+07f; hx = (int32_t)asuint(x);}  /* only positive handled */
  k += (hx>>23)-127;
  int32_t i = ((uint32_t)k&0x80000000)>>31;
  hx = (hx&0x007fffff)|((0x7f-i)<<23);
  float y = (float)(k+i);
  x = asfloat((uint32_t)hx);
  const float ivln10 = 4.3429449201e-01f, log10_2hi = 3.0102920532e-01f, log10_2lo = 7.9034151668e-07f;
#ifdef USE_FMA2
  float z = fmaf(y, log10_2lo, ivln10*my_logf(x));
  return fmaf(y, log10_2hi, z);
#else
  float z = y*log10_2lo + ivln10*my_logf(x);
  return z + y*log10_2hi;
#endif
}
int main(int argc, char **argv){ long dl=0, d10=0, n=0; uint32_t step = argc > 1 ? (uint32_t)atoi(argv[1]) : 1;
  for (uint64_t u = 1; u < 0x7f800000u; u += step) { float x = asfloat((uint32_t)u);
    if (u >= 0x00800000u && my_logf(x) != logf(x)) dl++;
    if (my_log10f(x) != log10f(x)) d10++; n++; }
  printf("n=%ld logf diff=%ld log10f diff=%ld\n", n, dl, d10); return 0; }
