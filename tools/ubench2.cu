// Gather micro-benchmarks (development tool): what does one warp-wide position gather cost on B200, as a function of
// how the 32 lanes' addresses fall into 32-byte sectors and 128-byte lines?  Decides between list layouts.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench2 tools/ubench2.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ double4 ld256(const double4 *p)
{
  double4 r;
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  return r;
}

// every thread walks its own index list (coalesced rows idx[k*nthreads + t]); PF gathers in flight
template <int MODE>   // 0: LDG.256 of double4; 1: two LDG.128; 2: three LDG.64 from SoA
__global__ void __launch_bounds__(128) k_gather(const double4 *__restrict__ pos, const double *__restrict__ sx, const int *__restrict__ idx, int nthreads, int nk, double *out)
{
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  double acc = 0;
  for (int k0 = 0; k0 < nk; k0 += 4) {
    int j[4];
#pragma unroll
    for (int u = 0; u < 4; u++) j[u] = __ldcs(idx + (size_t) (k0 + u) * nthreads + t);
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (MODE == 0) { double4 p = ld256(pos + j[u]); acc += p.x + p.y + p.z; }
      else if (MODE == 1) {
        const double2 *q = reinterpret_cast<const double2 *>(pos + j[u]);
        double2 a = __ldg(q), b = __ldg(q + 1); acc += a.x + a.y + b.x;
      } else { acc += __ldg(sx + j[u]) + __ldg(sx + (1 << 22) + j[u]) + __ldg(sx + (2 << 22) + j[u]); }
    }
  }
  out[t] = acc;
}

// shared-memory variants: SoA LDS.64 x3, and ATOMS.ADD.64 x3 on integer accumulators
template <int MODE>   // 0: 3 LDS.64; 1: 3 LDS.64 + 3 atomicAdd(u64); 2: LDS.128 {x,y} + LDS.64 z
__global__ void __launch_bounds__(128) k_smem(const unsigned short *__restrict__ slots, int nthreads, int nk, int iters, double *out)
{
  extern __shared__ double sm[];
  const int NS = 2048;
  double *s_x = sm, *s_y = sm + NS, *s_z = sm + 2 * NS;
  unsigned long long *f = reinterpret_cast<unsigned long long *>(sm + 3 * NS);
  for (int i = threadIdx.x; i < NS * (MODE == 1 ? 6 : 3); i += blockDim.x) sm[i] = i * 1e-3;
  __syncthreads();
  int t = (blockIdx.x * blockDim.x + threadIdx.x) % nthreads;
  double acc = 0;
  for (int it = 0; it < iters; it++)
    for (int k0 = 0; k0 < nk; k0 += 4) {
      int s[4];
#pragma unroll
      for (int u = 0; u < 4; u++) s[u] = slots[(size_t) (k0 + u) * nthreads + t];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        double a, b, c;
        if (MODE == 2) { double2 v = reinterpret_cast<double2 *>(sm)[s[u]]; a = v.x; b = v.y; c = s_z[s[u]]; }
        else { a = s_x[s[u]]; b = s_y[s[u]]; c = s_z[s[u]]; }
        acc += a + b + c;
        if (MODE == 1) {
          atomicAdd(&f[s[u]], (unsigned long long) __double_as_longlong(a));
          atomicAdd(&f[NS + s[u]], (unsigned long long) __double_as_longlong(b));
          atomicAdd(&f[2 * NS + s[u]], (unsigned long long) __double_as_longlong(c));
        }
      }
    }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (MODE == 1 ? (double) f[threadIdx.x] : 0.0);
}

static unsigned rng_state = 12345u;
static unsigned rnd() { rng_state = rng_state * 1664525u + 1013904223u; return rng_state >> 8; }

int main()
{
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  const int NP = 1 << 22;          // 4 M positions = 128 MB (L2-resident mostly, like a tile neighbourhood is L1/L2 resident)
  const int NT = 148 * 16 * 128;   // threads
  const int NK = 64;
  double4 *pos; double *sx, *out; int *idx; unsigned short *slots;
  CK(cudaMalloc(&pos, sizeof(double4) * NP)); CK(cudaMalloc(&sx, sizeof(double) * 3 * (1 << 22)));
  CK(cudaMemset(pos, 0, sizeof(double4) * NP)); CK(cudaMemset(sx, 0, sizeof(double) * 3 * (1 << 22)));
  CK(cudaMalloc(&out, sizeof(double) * NT)); CK(cudaMalloc(&idx, sizeof(int) * (size_t) NT * NK));
  CK(cudaMalloc(&slots, 2 * (size_t) NT * NK));
  int *h = (int *) malloc(sizeof(int) * (size_t) NT * NK);
  unsigned short *hs = (unsigned short *) malloc(2 * (size_t) NT * NK);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  // patterns: each warp works in a private window of W atoms around base (like a tile neighbourhood of ~500 atoms)
  struct Pat { const char *name; int share; int perline; int window; };
  const Pat pats[] = {
      {"32 lanes -> 32 random atoms in a 512-atom window (the md gather)", 1, 0, 512},
      {"pairs of lanes share an atom: 16 distinct sectors", 2, 0, 512},
      {"4 lanes share an atom: 8 distinct sectors", 4, 0, 512},
      {"32 distinct atoms, 4 consecutive atoms (one 128-byte line) per 4 lanes: 8 lines", 1, 4, 512},
      {"32 distinct atoms, 2 consecutive atoms per 2 lanes: 16 lines", 1, 2, 512},
      {"32 consecutive atoms (coalesced, 8 lines)", 1, 32, 512},
      {"32 random atoms in a 64-atom window (16 lines exist)", 1, 0, 64},
  };
  for (const Pat &pt : pats) {
    for (int w = 0; w < NT / 32; w++) {
      const int base = (int) (((size_t) w * 997) % (NP - 4096));
      for (int k = 0; k < NK; k++) {
        int grp[32];
        for (int l = 0; l < 32; l++) {
          int a;
          if (pt.perline) {
            if (l % pt.perline == 0) grp[l] = (int) (rnd() % (pt.window / pt.perline)) * pt.perline;
            else grp[l] = grp[l - 1] + 1;
            a = grp[l];
          } else {
            if (l % pt.share == 0) grp[l] = (int) (rnd() % pt.window);
            else grp[l] = grp[l - 1];
            a = grp[l];
          }
          h[(size_t) k * NT + w * 32 + l] = base + a;
        }
      }
    }
    CK(cudaMemcpy(idx, h, sizeof(int) * (size_t) NT * NK, cudaMemcpyHostToDevice));
    for (int mode = 0; mode < 3; mode++) {
      float best = 1e30f;
      for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        if (mode == 0) k_gather<0><<<NT / 128, 128>>>(pos, sx, idx, NT, NK, out);
        if (mode == 1) k_gather<1><<<NT / 128, 128>>>(pos, sx, idx, NT, NK, out);
        if (mode == 2) k_gather<2><<<NT / 128, 128>>>(pos, sx, idx, NT, NK, out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
      }
      CK(cudaGetLastError());
      const double warpsteps = (double) NT / 32 * NK;
      printf("%-82s %s  %7.3f ms  %6.2f SM-cycles per warp gather (at 1.9 GHz)\n", pt.name,
             mode == 0 ? "LDG.256  " : mode == 1 ? "2xLDG.128" : "3xLDG.64 ", best, best * 1e-3 * 1.9e9 * nsm / warpsteps);
    }
  }
  // shared memory
  struct SPat { const char *name; int kind; };
  const SPat spats[] = {{"smem: 32 lanes random slots in a 512-slot window", 0},
                        {"smem: per half-warp 16 ascending slots, gaps 1..12 (one atom's sorted list)", 1},
                        {"smem: per half-warp 16 consecutive slots", 2}};
  CK(cudaFuncSetAttribute(k_smem<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CK(cudaFuncSetAttribute(k_smem<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CK(cudaFuncSetAttribute(k_smem<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  const int SNT = 148 * 8 * 128;
  for (const SPat &sp : spats) {
    for (int w = 0; w < SNT / 32; w++)
      for (int k = 0; k < NK; k++) {
        int base = (int) (rnd() % 1400);
        int run = base;
        for (int l = 0; l < 32; l++) {
          int s;
          if (sp.kind == 0) s = base + (int) (rnd() % 512);
          else if (sp.kind == 1) { if (l % 16 == 0) run = base + (int) (rnd() % 64); run += 1 + (int) (rnd() % 12); s = run; }
          else { if (l % 16 == 0) run = base + (int) (rnd() % 64); run += 1; s = run; }
          hs[(size_t) k * SNT + w * 32 + l] = (unsigned short) (s & 2047);
        }
      }
    CK(cudaMemcpy(slots, hs, 2 * (size_t) SNT * NK, cudaMemcpyHostToDevice));
    for (int mode = 0; mode < 3; mode++) {
      const size_t sh = sizeof(double) * 2048 * (mode == 1 ? 6 : 3);
      float best = 1e30f;
      const int iters = 8;
      for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        if (mode == 0) k_smem<0><<<SNT / 128, 128, sh>>>(slots, SNT, NK, iters, out);
        if (mode == 1) k_smem<1><<<SNT / 128, 128, sh>>>(slots, SNT, NK, iters, out);
        if (mode == 2) k_smem<2><<<SNT / 128, 128, sh>>>(slots, SNT, NK, iters, out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
      }
      CK(cudaGetLastError());
      const double warpsteps = (double) SNT / 32 * NK * iters;
      printf("%-82s %s  %7.3f ms  %6.2f SM-cycles per warp gather (at 1.9 GHz)\n", sp.name,
             mode == 0 ? "3xLDS.64        " : mode == 1 ? "3xLDS.64+3xATOMS" : "LDS.128+LDS.64  ", best,
             best * 1e-3 * 1.9e9 * nsm / warpsteps);
    }
  }
  return 0;
}
