// noalloc.cu -- b200md_ke_noalloc: the kinetic-energy sum of b200md_thermo (thermo_ke in nve.cu: k_ke + the fixed-order
// reduction of the per-block partials; ComputeTemp::compute_scalar, src/compute_temp.cpp:72-100) for callers that may
// not allocate device memory (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in
// the reference.
//
// The per-block partials live in a file-scope __device__ array of fixed size instead of the context's growable
// `partials` buffer.  Every GPU that loads the library holds the array (512 KiB) until the process exits.  Every call
// on a device shares it: calls that overlap on the GPU (different streams, different host threads) race.  Nothing in
// it has to start at zero -- each call writes partial b of every block b < nblocks before the reduction reads exactly
// those -- so the contents a previous call left behind are never read.
//
// In host code `g_ke_scratch` is only a stand-in for the symbol, not a device address; the address the kernels are
// given comes from cudaGetSymbolAddress, taken once per device with that device current.
#include <mutex>

#include "ctx.cuh"

#define KE_NOALLOC_MAX_BLOCKS 65536   // x 256 atoms per block = B200MD_KE_NOALLOC_MAX_ATOMS (16 777 216)
#define KE_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double g_ke_scratch[KE_NOALLOC_MAX_BLOCKS];

// same block shape and summation order as k_ke (nve.cu), the type read from an int array instead of pos.w
__global__ void __launch_bounds__(256)
k_ke_noalloc(const double *__restrict__ vx, const double *__restrict__ vy, const double *__restrict__ vz,
             const int *__restrict__ type, const double *__restrict__ mass, int n, double *__restrict__ partials)
{
  __shared__ double s_red[8];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double t = 0.0;
  if (i < n) {
    double a = vx[i], b = vy[i], cz = vz[i];
    t = (a * a + b * b + cz * cz) * mass[type[i]];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int k = 0; k < 8; k++) s += s_red[k];
    partials[blockIdx.x] = s;
  }
}

// fixed-order sum of the partials, as k_reduce_partials (pair_lj.cu) for one component
__global__ void __launch_bounds__(1024) k_ke_noalloc_reduce(const double *__restrict__ partials, int nblocks, double *__restrict__ out)
{
  __shared__ double s[1024];
  double t = 0.0;
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x) t += partials[b];
  s[threadIdx.x] = t;
  __syncthreads();
  for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = s[0];
}

static std::mutex g_ke_mutex;
static double *g_ke_scratch_addr[KE_NOALLOC_MAX_DEVICES];   // device address of g_ke_scratch, per device; set once

extern "C" int b200md_ke_noalloc_max_atoms(void) { return KE_NOALLOC_MAX_BLOCKS * 256; }

extern "C" int b200md_ke_noalloc(int device, int n, const double *vx, const double *vy, const double *vz, const int *type,
                                 const double *mass, double *sum_mv2, void *stream)
{
  b200md_ctx *c = nullptr;   // errors go to the library-wide string, b200md_last_error(NULL)
  if (device < 0 || device >= KE_NOALLOC_MAX_DEVICES)
    return b200md_fail(c, B200MD_ERR_ARG, "b200md_ke_noalloc: device %d outside [0, %d)", device, KE_NOALLOC_MAX_DEVICES);
  if (n < 0 || n > KE_NOALLOC_MAX_BLOCKS * 256)
    return b200md_fail(c, B200MD_ERR_ARG,
                       "b200md_ke_noalloc: %d atoms exceed the fixed scratch array (at most %d atoms = %d blocks of 256)", n,
                       KE_NOALLOC_MAX_BLOCKS * 256, KE_NOALLOC_MAX_BLOCKS);
  if (!sum_mv2 || (n > 0 && (!vx || !vy || !vz || !type || !mass)))
    return b200md_fail(c, B200MD_ERR_ARG, "b200md_ke_noalloc: NULL array");
  int prev = -1;
  CU(cudaGetDevice(&prev));
  if (prev != device) CU(cudaSetDevice(device));
  double *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_ke_mutex);
    if (!g_ke_scratch_addr[device]) {
      void *p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_ke_scratch);
      if (e != cudaSuccess || !p) {
        if (prev != device) cudaSetDevice(prev);
        return b200md_fail(c, B200MD_ERR_CUDA, "b200md_ke_noalloc: cudaGetSymbolAddress(g_ke_scratch) on device %d: %s", device,
                           cudaGetErrorString(e));
      }
      g_ke_scratch_addr[device] = static_cast<double *>(p);
    }
    scratch = g_ke_scratch_addr[device];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = (n + 255) / 256;
  if (blocks > 0) k_ke_noalloc<<<blocks, 256, 0, st>>>(vx, vy, vz, type, mass, n, scratch);
  k_ke_noalloc_reduce<<<1, 1024, 0, st>>>(scratch, blocks, sum_mv2);
  cudaError_t e = cudaGetLastError();
  if (prev != device) cudaSetDevice(prev);
  if (e != cudaSuccess) return b200md_fail(c, B200MD_ERR_CUDA, "b200md_ke_noalloc: kernel launch: %s", cudaGetErrorString(e));
  return B200MD_OK;
}
