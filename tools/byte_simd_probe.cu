#include <cstdint>
extern "C" __global__ void k_max(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vmaxu4(a[i], b[i]); }
extern "C" __global__ void k_min(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vminu4(a[i], b[i]); }
extern "C" __global__ void k_cmpgt(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vcmpgtu4(a[i], b[i]); }
extern "C" __global__ void k_subus(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vsubus4(a[i], b[i]); }
extern "C" __global__ void k_addus(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vaddus4(a[i], b[i]); }
extern "C" __global__ void k_absd(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vabsdiffu4(a[i], b[i]); }
extern "C" __global__ void k_setgt(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vsetgtu4(a[i], b[i]); }
extern "C" __global__ void k_max2(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vmaxu2(a[i], b[i]); }
extern "C" __global__ void k_cmpgt2(const uint32_t* a, const uint32_t* b, uint32_t* o) { int i = threadIdx.x; o[i] = __vcmpgtu2(a[i], b[i]); }
