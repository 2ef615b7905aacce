// TEST INFRASTRUCTURE -- never shipped, never loaded by the package.
//
// A lockstep CUDA emulation shim: tests/cudaemu/build.py rewrites the `<<<...>>>` launches of
// lammps-fugaku-threadpool_b200/csrc/*.cu into emu::launch(...) calls and compiles the SAME kernel and host sources
// with g++ against this header into tests/cudaemu/_build/libb200md_emu.so.  Purpose: desk-check kernel and host
// LOGIC (indexing, flag protocols, list layouts, call order, buffer sizes) in a container without a GPU, before the
// GPU tests run on the box.  It says nothing about performance and is not bit-identical to the device (no FMA
// contraction here, exact reciprocal instead of the MUFU seed); parity claims rest on the -m gpu tests only.
//
// Execution model: per host thread (= per emulated rank) one scheduler; every CUDA thread of a block is a fiber (emu.cpp); blocks run one after the
// other.  __syncthreads and the *_sync warp primitives are real rendezvous between fibers, exited threads count as
// arrived.  "Device" memory is host memory between PROT_NONE guard pages (end-aligned), so an out-of-bounds device
// access faults instead of passing silently.  Streams and events are no-ops: every call completes before it returns.
#pragma once
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#ifndef __align__
#define __align__(n) __attribute__((aligned(n)))
#endif
#define __shared__ static thread_local   /* one block at a time per host thread (= per emulated rank) */
#ifndef __restrict__
#define __restrict__ __restrict
#endif

struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { return {x, y}; }
struct double3 { double x, y, z; };
struct double4 { double x, y, z, w; };
static inline double3 make_double3(double x, double y, double z) { return {x, y, z}; }
static inline double4 make_double4(double x, double y, double z, double w) { return {x, y, z, w}; }
static inline double2 make_double2(double x, double y) { return {x, y}; }
static inline int4 make_int4(int x, int y, int z, int w) { return {x, y, z, w}; }
static inline int2 make_int2(int x, int y) { return {x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }

// ---- runtime API -------------------------------------------------------------------------------------
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
typedef struct emu_stream *cudaStream_t;
typedef struct emu_event *cudaEvent_t;
enum { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
struct cudaDeviceProp { int major, minor, multiProcessorCount; char name[64]; };

namespace emu {
void *dev_alloc(size_t bytes);
void dev_free(void *p);
// one block at a time; body() is the kernel call with its arguments (evaluated inside each thread's fiber)
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
void block_barrier();
void warp_barrier();
struct ThreadState { uint3 threadIdx, blockIdx; dim3 blockDim, gridDim; int lane, warp; };
extern thread_local ThreadState *cur;
extern unsigned long long *warp_scratch();   // [32] of the running thread's warp
extern unsigned warp_livemask();             // lanes of the running thread's warp that have not exited
extern thread_local void *dyn_smem;
extern thread_local int launches;
}    // namespace emu

#define threadIdx (emu::cur->threadIdx)
#define blockIdx (emu::cur->blockIdx)
#define blockDim (emu::cur->blockDim)
#define gridDim (emu::cur->gridDim)

static inline const char *cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int *n) { *n = 64; return cudaSuccess; }   // one emulated device per host thread
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
// a __device__ variable is a plain host variable here: its address is the "device" address
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, const T &sym) { *p = (void *) &sym; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int)
{
  memset(p, 0, sizeof(*p));
  p->major = 10;
  p->multiProcessorCount = 148;
  strcpy(p->name, "cudaemu (host fibers)");
  return cudaSuccess;
}
template <class T> static inline cudaError_t cudaMalloc(T **p, size_t bytes)
{
  *p = (T *) emu::dev_alloc(bytes);
  return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
static inline cudaError_t cudaFree(void *p) { emu::dev_free(p); return cudaSuccess; }
template <class T> static inline cudaError_t cudaMallocHost(T **p, size_t bytes) { *p = (T *) malloc(bytes ? bytes : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
static inline cudaError_t cudaFreeHost(void *p) { free(p); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { if (n) memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { if (n) memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = nullptr) { if (n) memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = (cudaStream_t) malloc(8); return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = (cudaEvent_t) malloc(8); return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { *e = (cudaEvent_t) malloc(8); return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { free(e); return cudaSuccess; }
// everything is synchronous, so an event is just the wall-clock time of its record (malloc(8) above holds it)
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  *(double *) e = ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
  return cudaSuccess;
}
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float) (*(double *) b - *(double *) a); return cudaSuccess; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }

// ---- device intrinsics ---------------------------------------------------------------------------------
static inline void __syncthreads() { emu::block_barrier(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::warp_barrier(); }
template <class T> static inline T __ldg(const T *p) { return *p; }
template <class T> static inline T __ldcs(const T *p) { return *p; }
template <class T> static inline T __ldcg(const T *p) { return *p; }
template <class T> static inline void __stcs(T *p, T v) { *p = v; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __fma_rn(double a, double b, double c) { return fma(a, b, c); }
static inline int __double2loint(double d) { uint64_t u; memcpy(&u, &d, 8); return (int) (u & 0xffffffffu); }
static inline int __double2hiint(double d) { uint64_t u; memcpy(&u, &d, 8); return (int) (u >> 32); }
static inline double __hiloint2double(int hi, int lo) { uint64_t u = ((uint64_t) (uint32_t) hi << 32) | (uint32_t) lo; double d; memcpy(&d, &u, 8); return d; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline long long __double_as_longlong(double d) { long long i; memcpy(&i, &d, 8); return i; }
static inline double __longlong_as_double(long long i) { double d; memcpy(&d, &i, 8); return d; }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned) v) : 32; }
using std::max;
using std::min;
static inline double rsqrt(double x) { return 1.0 / sqrt(x); }

// ranks are threads of this process: a "peer mapping" is the pointer itself, a system-scope fence a host fence
struct cudaIpcMemHandle_t { char reserved[64]; };
enum { cudaIpcMemLazyEnablePeerAccess = 1 };
static inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *h, void *p) { memset(h, 0, sizeof(*h)); memcpy(h->reserved, &p, sizeof(p)); return cudaSuccess; }
static inline cudaError_t cudaIpcOpenMemHandle(void **p, cudaIpcMemHandle_t h, unsigned) { memcpy(p, h.reserved, sizeof(*p)); return cudaSuccess; }
static inline cudaError_t cudaIpcCloseMemHandle(void *) { return cudaSuccess; }
static inline void __threadfence_system() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline long long clock64()
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long) ts.tv_sec * 1000000000LL + ts.tv_nsec;   // "cycles" at 1 GHz
}

// one fiber at a time: plain read-modify-write is atomic here
template <class T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicMax(T *p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T> static inline T atomicMin(T *p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T> static inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; return o; }
template <class T> static inline T atomicExch(T *p, T v) { T o = *p; *p = v; return o; }
template <class T> static inline T atomicCAS(T *p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }

// warp collectives: deposit, rendezvous, read, rendezvous.  Every mask is taken as "all lanes of the warp that are
// still running" (how the kernels use them); a lane that has exited contributes nothing.
static inline unsigned __activemask() { return emu::warp_livemask(); }
template <class T> static inline unsigned long long emu_bits(T v) { unsigned long long u = 0; static_assert(sizeof(T) <= 8, "8-byte shuffles at most"); memcpy(&u, &v, sizeof(T)); return u; }
template <class T> static inline T emu_unbits(unsigned long long u) { T v; memcpy(&v, &u, sizeof(T)); return v; }
template <class T> static inline T emu_shfl_from(T v, int src)
{
  unsigned long long *s = emu::warp_scratch();
  s[emu::cur->lane] = emu_bits(v);
  emu::warp_barrier();
  T r = v;
  if (src >= 0 && src < 32 && (emu::warp_livemask() >> src & 1u)) r = emu_unbits<T>(s[src]);
  emu::warp_barrier();
  return r;
}
template <class T> static inline T __shfl_sync(unsigned, T v, int src, int width = 32) { return emu_shfl_from(v, (emu::cur->lane & ~(width - 1)) | (src & (width - 1))); }
template <class T> static inline T __shfl_down_sync(unsigned, T v, unsigned d, int width = 32)
{
  int src = emu::cur->lane + (int) d;
  if ((src & ~(width - 1)) != (emu::cur->lane & ~(width - 1))) src = -1;
  return emu_shfl_from(v, src);
}
template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d, int width = 32)
{
  int src = emu::cur->lane - (int) d;
  if (src < (emu::cur->lane & ~(width - 1))) src = -1;
  return emu_shfl_from(v, src);
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m, int width = 32) { return emu_shfl_from(v, emu::cur->lane ^ m); }
static inline unsigned __ballot_sync(unsigned, int pred)
{
  unsigned long long *s = emu::warp_scratch();
  s[emu::cur->lane] = pred ? 1 : 0;
  emu::warp_barrier();
  unsigned live = emu::warp_livemask(), r = 0;
  for (int l = 0; l < 32; l++)
    if ((live >> l & 1u) && s[l]) r |= 1u << l;
  emu::warp_barrier();
  return r;
}
// packed FP32 pair arithmetic of sm_100 (crt/sm_100_rt.h): two independent IEEE operations
static inline float2 __fadd2_rn(float2 a, float2 b) { return {a.x + b.x, a.y + b.y}; }
static inline float2 __fmul2_rn(float2 a, float2 b) { return {a.x * b.x, a.y * b.y}; }
static inline float2 __ffma2_rn(float2 a, float2 b, float2 c) { return {fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)}; }
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, !pred) == 0; }
static inline int __reduce_max_sync(unsigned, int v)
{
  unsigned long long *s = emu::warp_scratch();
  s[emu::cur->lane] = emu_bits(v);
  emu::warp_barrier();
  unsigned live = emu::warp_livemask();
  int r = v;
  for (int l = 0; l < 32; l++)
    if (live >> l & 1u) r = std::max(r, emu_unbits<int>(s[l]));
  emu::warp_barrier();
  return r;
}
static inline int __reduce_add_sync(unsigned, int v)
{
  unsigned long long *s = emu::warp_scratch();
  s[emu::cur->lane] = emu_bits(v);
  emu::warp_barrier();
  unsigned live = emu::warp_livemask();
  int r = 0;
  for (int l = 0; l < 32; l++)
    if (live >> l & 1u) r += emu_unbits<int>(s[l]);
  emu::warp_barrier();
  return r;
}
