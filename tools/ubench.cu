// Micro-benchmarks that decide the pair-kernel design on B200 (not product code):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu && tools/ubench
// Each line: what, total lane-ops, ms, lane-ops per clock per SM at the clock the run itself measured.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void k_clock(long long *out, int iters) {
    long long t0 = clock64();
    float a = threadIdx.x;
    for (int i = 0; i < iters; i++) a = a * 1.0001f + 0.5f;
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = (long long)a; }
}

// independent DFMA chains per thread
template <int ILP>
__global__ void k_dfma(double *out, int iters, double s) {
    double a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; k++) a[k] = threadIdx.x + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < ILP; k++) a[k] = fma(a[k], s, 0.25);
    }
    double r = 0;
#pragma unroll
    for (int k = 0; k < ILP; k++) r += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// DFMA with a fraction of lanes active (does a half-empty warp cost a full DP issue?)
__global__ void k_dfma_masked(double *out, int iters, double s, int active) {
    double a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    if ((threadIdx.x & 31) < active) {
        for (int i = 0; i < iters; i++) {
            a0 = fma(a0, s, 0.25); a1 = fma(a1, s, 0.25); a2 = fma(a2, s, 0.25); a3 = fma(a3, s, 0.25);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}

// reciprocal: 1.0/x as the compiler emits it
__global__ void k_drcp(double *out, int iters, double s) {
    double a0 = 1.0 + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    for (int i = 0; i < iters; i++) {
        a0 = 1.0 / (a0 + s); a1 = 1.0 / (a1 + s); a2 = 1.0 / (a2 + s); a3 = 1.0 / (a3 + s);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}

// shared-memory gather: LDS.64 x3 from random slots (x[] y[] z[] layout), slots from a pre-generated list in registers
template <int MODE>   // 0: three LDS.64 SoA; 1: one LDS.128 (packed 16-byte record); 2: SoA + atomicAdd(double) x3 back
__global__ void k_smem_gather(double *out, const unsigned short *slots, int nslots, int iters, int nstaged) {
    extern __shared__ double sm[];
    double *sx = sm, *sy = sm + nstaged, *sz = sm + 2 * nstaged;
    double *fx = sm + 3 * nstaged, *fy = sm + 4 * nstaged, *fz = sm + 5 * nstaged;
    for (int i = threadIdx.x; i < nstaged * (MODE == 2 ? 6 : 3); i += blockDim.x) sm[i] = i * 0.001;
    __syncthreads();
    double acc = 0;
    const unsigned short *my = slots + (size_t)(blockIdx.x * blockDim.x + threadIdx.x) % 4096;
    for (int it = 0; it < iters; it++) {
#pragma unroll 4
        for (int k = 0; k < nslots; k++) {
            int s = my[k * 4096] % nstaged;     // coalesced 16-bit list row
            if (MODE == 0) {
                acc += sx[s] + sy[s] + sz[s];
            } else if (MODE == 1) {
                double2 v = reinterpret_cast<double2 *>(sm)[s];
                acc += v.x + v.y;
            } else {
                double a = sx[s], b = sy[s], c = sz[s];
                acc += a + b + c;
                atomicAdd(&fx[s], a); atomicAdd(&fy[s], b); atomicAdd(&fz[s], c);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (MODE == 2 ? fx[threadIdx.x % nstaged] : 0.0);
}

// global red.add.f64: each thread adds to nadd addresses near its own index (spatial locality like an MD scatter)
__global__ void k_red(double *f, const int *idx, int n, int nadd) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int k = 0; k < nadd; k++) {
        int j = idx[(size_t)k * n + i];
        atomicAdd(&f[j], 1.0);
    }
}
__global__ void k_mkidx(int *idx, int n, int nadd, int spread) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned h = i * 2654435761u;
    for (int k = 0; k < nadd; k++) {
        h = h * 1664525u + 1013904223u;
        int j = i + (int)((h >> 8) % (2 * spread + 1)) - spread;
        j = j < 0 ? 0 : (j >= n ? n - 1 : j);
        idx[(size_t)k * n + i] = j;
    }
}

// indexed shuffle gather of a double3 (6 SHFL) per step
__global__ void k_shfl(double *out, int iters) {
    double x = threadIdx.x, y = x + 1, z = x + 2, acc = 0;
    int src = (threadIdx.x * 7 + 3) & 31;
    for (int i = 0; i < iters; i++) {
        acc += __shfl_sync(0xffffffffu, x, src) + __shfl_sync(0xffffffffu, y, src) + __shfl_sync(0xffffffffu, z, src);
        src = (src * 5 + 1) & 31;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

static float timeit(void (*launch)(void *), void *arg, int reps = 3) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(arg); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0)); launch(arg); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

struct A { double *out; int iters; unsigned short *slots; int *idx; double *f; int n; int mode; int nstaged; int active; };
static const int NB = 148 * 8, NT = 256;

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    printf("device %s, %d SMs\n", p.name, nsm);
    A a{}; CK(cudaMalloc(&a.out, sizeof(double) * NB * NT * 2));
    // clock: measure cycles of a fixed kernel against event time
    long long *clk; CK(cudaMalloc(&clk, 16));
    double mhz = 0;
    {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k_clock<<<1, 32>>>(clk, 1 << 22); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_clock<<<1, 32>>>(clk, 1 << 24); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); long long h[2]; cudaMemcpy(h, clk, 16, cudaMemcpyDeviceToHost);
        mhz = h[0] / (ms * 1e3); printf("idle-ish SM clock %.0f MHz\n", mhz);
    }
    auto report = [&](const char *what, double laneops, float ms) {
        printf("%-58s %10.3f ms  %8.2f lane-ops/clk/SM (at %.0f MHz)  %8.2f Gops/s\n", what, ms,
               laneops / (ms * 1e-3) / (mhz * 1e6) / nsm, mhz, laneops / (ms * 1e-3) / 1e9);
    };
    a.iters = 4096;
    report("DFMA ILP1 (8 warps x 8 blocks/SM)", (double)NB * NT * a.iters * 1,
           timeit([](void *v) { A *a = (A *)v; k_dfma<1><<<NB, NT>>>(a->out, a->iters, 0.999); }, &a));
    report("DFMA ILP4", (double)NB * NT * a.iters * 4,
           timeit([](void *v) { A *a = (A *)v; k_dfma<4><<<NB, NT>>>(a->out, a->iters, 0.999); }, &a));
    report("DFMA ILP8", (double)NB * NT * a.iters * 8,
           timeit([](void *v) { A *a = (A *)v; k_dfma<8><<<NB, NT>>>(a->out, a->iters, 0.999); }, &a));
    for (int act : {32, 24, 16, 8, 1}) {
        a.active = act; char nm[80]; snprintf(nm, 80, "DFMA ILP4, %d of 32 lanes active (ops counted as 32)", act);
        report(nm, (double)NB * NT * a.iters * 4,
               timeit([](void *v) { A *a = (A *)v; k_dfma_masked<<<NB, NT>>>(a->out, a->iters, 0.999, a->active); }, &a));
    }
    a.iters = 1024;
    report("1.0/x double (compiler sequence), per reciprocal", (double)NB * NT * a.iters * 4,
           timeit([](void *v) { A *a = (A *)v; k_drcp<<<NB, NT>>>(a->out, a->iters, 0.5); }, &a));
    report("indexed SHFL gather of a double3 (6 SHFL), per gather", (double)NB * NT * 4096,
           timeit([](void *v) { A *a = (A *)v; k_shfl<<<NB, NT>>>(a->out, 4096); }, &a));
    // shared gathers
    {
        int nslots = 64; unsigned short *h = (unsigned short *)malloc(2 * 4096 * nslots);
        // sorted-ish lists: neighbouring threads read nearby slots (like cell-ordered atoms)
        for (int k = 0; k < nslots; k++) for (int t = 0; t < 4096; t++) {
            unsigned r = (unsigned)rand();
            h[k * 4096 + t] = (unsigned short)((k * 20 + (t % 128) / 4 + r % 40) % 1300);
        }
        CK(cudaMalloc(&a.slots, 2 * 4096 * nslots)); CK(cudaMemcpy(a.slots, h, 2 * 4096 * nslots, cudaMemcpyHostToDevice));
        a.nstaged = 1300; a.iters = 8; int nb = 148 * 4;
        for (int mode = 0; mode < 3; mode++) {
            a.mode = mode;
            size_t sh = sizeof(double) * a.nstaged * (mode == 2 ? 6 : 3);
            auto L = [](void *v) {
                A *a = (A *)v; size_t sh = sizeof(double) * a->nstaged * (a->mode == 2 ? 6 : 3); int nb = 148 * 4;
                if (a->mode == 0) k_smem_gather<0><<<nb, 128, sh>>>(a->out, a->slots, 64, a->iters, a->nstaged);
                if (a->mode == 1) k_smem_gather<1><<<nb, 128, sh>>>(a->out, a->slots, 64, a->iters, a->nstaged);
                if (a->mode == 2) k_smem_gather<2><<<nb, 128, sh>>>(a->out, a->slots, 64, a->iters, a->nstaged);
            };
            CK(cudaFuncSetAttribute(k_smem_gather<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
            (void)sh;
            const char *nm[3] = {"smem gather 3xLDS.64 (x[]y[]z[]), per neighbor", "smem gather 1xLDS.128 packed, per neighbor",
                                 "smem gather + 3x atomicAdd(double) shared, per neighbor"};
            report(nm[mode], (double)nb * 128 * a.iters * nslots, timeit(L, &a));
        }
    }
    // global red.add.f64
    {
        a.n = 4 << 20; int nadd = 12;
        CK(cudaMalloc(&a.f, sizeof(double) * a.n)); CK(cudaMalloc(&a.idx, sizeof(int) * (size_t)a.n * nadd));
        for (int spread : {32, 1024, 1 << 20}) {
            k_mkidx<<<(a.n + 255) / 256, 256>>>(a.idx, a.n, nadd, spread); CK(cudaDeviceSynchronize());
            char nm[80]; snprintf(nm, 80, "global atomicAdd(double) no return, spread +-%d, per add", spread);
            report(nm, (double)a.n * nadd,
                   timeit([](void *v) { A *a = (A *)v; k_red<<<(a->n + 255) / 256, 256>>>(a->f, a->idx, a->n, 12); }, &a));
        }
    }
    return 0;
}
