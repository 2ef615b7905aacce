// capi.cu -- the C ABI of libb200md.so (include/b200md.h): context lifetime, settings, atom upload /
// download, pair-style tables, and the device-resident Verlet loop (src/verlet.cpp:93-162, 229-360).
// No CPU fallback: b200md_create fails with B200MD_ERR_CUDA when no usable sm_100 device exists.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <new>
#include <vector>

#include "ctx.cuh"

int nve_final_initial(b200md_ctx *c, double dtv, double dtf);
bool chain_post_usable(const b200md_ctx *c);
int chain_post_fused(b200md_ctx *c, int64_t ntimestep);
int comm_init(b200md_ctx *c, const void *nccl_id);
void comm_destroy(b200md_ctx *c);

static char g_err[512] = "no error";

int b200md_fail(b200md_ctx *c, int code, const char *fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(c ? c->err : g_err, 512, fmt, ap);
  va_end(ap);
  return code;
}

extern "C" const char *b200md_last_error(const b200md_ctx *c) { return c ? c->err : g_err; }

extern "C" int b200md_create(b200md_ctx **out, int device, int nranks, int rank, const void *nccl_id)
{
  b200md_ctx *c = nullptr;
  if (!out) return b200md_fail(c, B200MD_ERR_ARG, "b200md_create: out == NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return b200md_fail(c, B200MD_ERR_CUDA, "no CUDA device (%s); libb200md has no CPU fallback",
                       e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return b200md_fail(c, B200MD_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return b200md_fail(c, B200MD_ERR_CUDA, "cudaGetDeviceProperties failed");
  if (prop.major != 10)
    return b200md_fail(c, B200MD_ERR_CUDA, "device %d is sm_%d%d; libb200md is built for sm_100a only", device, prop.major, prop.minor);
  if (nranks < 1 || rank < 0 || rank >= nranks) return b200md_fail(c, B200MD_ERR_ARG, "bad nranks/rank %d/%d", nranks, rank);
  if (cudaSetDevice(device) != cudaSuccess) return b200md_fail(c, B200MD_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  c = new (std::nothrow) b200md_ctx();
  if (!c) return b200md_fail(c, B200MD_ERR_ARG, "out of host memory");
  c->device = device;
  c->nranks = nranks;
  c->rank = rank;
  for (int k = 0; k <= B200MD_MAXTYPES; k++) c->mass_h[k] = 0.0;
  strcpy(c->err, "no error");
  // defaults chosen from hardware measurements; b200md_tune changes them per context
  c->overlap = nranks > 1;           // halo on its own stream under the ghost-free tiles' force kernels
  c->host_pipeline = 1;
  c->debug_timing = getenv("B200MD_DEBUG_TIMING") != nullptr;   // diagnostic only, read once
  *out = c;
  CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CU(cudaMalloc((void **) &c->red_d, 32 * sizeof(double)));
  CU(cudaMemset(c->red_d, 0, 32 * sizeof(double)));
  CU(cudaMallocHost((void **) &c->red_h, 32 * sizeof(double)));
  CU(cudaMalloc((void **) &c->flags_d, 8 * sizeof(int)));
  CU(cudaMemset(c->flags_d, 0, 8 * sizeof(int)));
  CU(cudaMallocHost((void **) &c->flags_h, (size_t) (64 + 32 * (nranks + 1)) * sizeof(int)));
  memset(c->flags_h, 0, (size_t) (64 + 32 * (nranks + 1)) * sizeof(int));
  if (nranks > 1) TRY(comm_init(c, nccl_id));
  return B200MD_OK;
}

extern "C" int b200md_destroy(b200md_ctx *c)
{
  if (!c) return B200MD_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  comm_destroy(c);
  dev_free(c, c->mass_d); dev_free(c, c->pos); dev_free(c, c->pos_tmp); dev_free(c, c->xhold);
  for (int k = 0; k < 3; k++) { dev_free(c, c->v[k]); dev_free(c, c->v_tmp[k]); dev_free(c, c->f[k]); }
  dev_free(c, c->tag); dev_free(c, c->tag_tmp); dev_free(c, c->image); dev_free(c, c->image_tmp);
  dev_free(c, c->rho); dev_free(c, c->fp); dev_free(c, c->posfp); dev_free(c, c->cutneighsq_d);
  dev_free(c, c->bin_count); dev_free(c, c->bin_start); dev_free(c, c->atom_bin); dev_free(c, c->cell_list);
  dev_free(c, c->perm); dev_free(c, c->scan_tmp); dev_free(c, c->rec); dev_free(c, c->pbin_start); dev_free(c, c->atom_sub); dev_free(c, c->numneigh); dev_free(c, c->neigh); dev_free(c, c->numneigh_in); dev_free(c, c->neigh_in); dev_free(c, c->tile_ghost); dev_free(c, c->tile_list); dev_free(c, c->tile_pos);
  dev_free(c, c->send_idx); dev_free(c, c->border_blockcnt); dev_free(c, c->border_total); dev_free(c, c->send_dir);
  dev_free(c, c->sendbuf_i); dev_free(c, c->allcounts); dev_free(c, c->sendbuf4); dev_free(c, c->recvbuf4); dev_free(c, c->sendbuf1);
  dev_free(c, c->partials); dev_free(c, c->stage); dev_free(c, c->peratom); dev_free(c, c->special.n3); dev_free(c, c->special.tags);
  bond_release(c);
  if (c->lj.d) cudaFree(c->lj.d);
  if (c->eam.frho) cudaFree(c->eam.frho);
  if (c->eam.rhor) cudaFree(c->eam.rhor);
  if (c->eam.z2r) cudaFree(c->eam.z2r);
  if (c->eam.rhoA) cudaFree(c->eam.rhoA);
  if (c->eam.tab1) cudaFree(c->eam.tab1);
  if (c->eam.type2frho) cudaFree(c->eam.type2frho);
  if (c->eam.type2rhor) cudaFree(c->eam.type2rhor);
  if (c->eam.type2z2r) cudaFree(c->eam.type2z2r);
  if (c->red_d) cudaFree(c->red_d);
  if (c->red_h) cudaFreeHost(c->red_h);
  if (c->flags_d) cudaFree(c->flags_d);
  if (c->flags_h) cudaFreeHost(c->flags_h);
  for (auto &e : c->kev) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
  if (c->s_in) {
    cudaStreamDestroy(c->s_in);
    cudaStreamDestroy(c->s_out);
    for (auto &e : c->hs_ev) cudaEventDestroy(e);
  }
  if (c->s_comm) {
    cudaStreamDestroy(c->s_comm);
    cudaEventDestroy(c->ev_main);
    cudaEventDestroy(c->ev_halo);
  }
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return B200MD_OK;
}

extern "C" double b200md_memory_usage(const b200md_ctx *c) { return c ? c->bytes_held : 0.0; }

#define NEED_CTX()                                                                 \
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");             \
  if (cudaSetDevice(c->device) != cudaSuccess) return b200md_fail(c, B200MD_ERR_CUDA, "cudaSetDevice failed")

extern "C" int b200md_set_units(b200md_ctx *c, double boltz, double mvv2e, double ftm2v, double nktv2p)
{
  NEED_CTX();
  c->boltz = boltz; c->mvv2e = mvv2e; c->ftm2v = ftm2v; c->nktv2p = nktv2p;
  return B200MD_OK;
}

extern "C" int b200md_set_box(b200md_ctx *c, const double boxlo[3], const double boxhi[3], const int periodic[3],
                              const int procgrid[3])
{
  NEED_CTX();
  int np = procgrid[0] * procgrid[1] * procgrid[2];
  if (np != c->nranks) return b200md_fail(c, B200MD_ERR_ARG, "procgrid %dx%dx%d != nranks %d", procgrid[0], procgrid[1], procgrid[2], c->nranks);
  for (int k = 0; k < 3; k++) {
    if (!(boxhi[k] > boxlo[k])) return b200md_fail(c, B200MD_ERR_ARG, "box hi <= lo in dim %d", k);
    c->boxlo[k] = boxlo[k]; c->boxhi[k] = boxhi[k]; c->prd[k] = boxhi[k] - boxlo[k];
    c->periodic[k] = periodic[k]; c->procgrid[k] = procgrid[k];
  }
  c->myloc[0] = c->rank % procgrid[0];
  c->myloc[1] = (c->rank / procgrid[0]) % procgrid[1];
  c->myloc[2] = c->rank / (procgrid[0] * procgrid[1]);
  for (int k = 0; k < 3; k++) {
    // uniform cuts, as CommBrick::setup / Domain::set_local_box (src/domain.cpp: sublo = boxlo + prd*split)
    c->sublo[k] = c->boxlo[k] + c->prd[k] * ((double) c->myloc[k] / c->procgrid[k]);
    c->subhi[k] = (c->myloc[k] == c->procgrid[k] - 1) ? c->boxhi[k] : c->boxlo[k] + c->prd[k] * ((double) (c->myloc[k] + 1) / c->procgrid[k]);
  }
  c->have_box = true;
  c->list_valid = false;
  return B200MD_OK;
}

extern "C" int b200md_set_mass(b200md_ctx *c, int ntypes, const double *mass)
{
  NEED_CTX();
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES) return b200md_fail(c, B200MD_ERR_ARG, "ntypes %d out of range 1..%d", ntypes, B200MD_MAXTYPES);
  c->ntypes = ntypes;
  c->mass_h[0] = 1.0;
  for (int k = 1; k <= ntypes; k++) {
    if (!(mass[k] > 0.0)) return b200md_fail(c, B200MD_ERR_ARG, "mass of type %d not set", k);
    c->mass_h[k] = mass[k];
  }
  TRY(dev_ensure(c, c->mass_d, B200MD_MAXTYPES + 1));
  CU(cudaMemcpyAsync(c->mass_d.p, c->mass_h, (ntypes + 1) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

// ---- atoms -----------------------------------------------------------------------------------
__global__ void k_pack_pos(const double *__restrict__ x3, const int *__restrict__ type, double4 *__restrict__ pos, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  pos[i] = make_double4(x3[3 * i], x3[3 * i + 1], x3[3 * i + 2], type_as_w(type[i]));
}
__global__ void k_aos_to_soa(const double *__restrict__ a3, double *__restrict__ a, double *__restrict__ b, double *__restrict__ cc, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  a[i] = a3[3 * i]; b[i] = a3[3 * i + 1]; cc[i] = a3[3 * i + 2];
}
__global__ void k_soa_to_aos(const double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ cc, double *__restrict__ a3, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  a3[3 * i] = a[i]; a3[3 * i + 1] = b[i]; a3[3 * i + 2] = cc[i];
}
__global__ void k_unpack_pos(const double4 *__restrict__ pos, double *__restrict__ x3, int *__restrict__ type, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  x3[3 * i] = p.x; x3[3 * i + 1] = p.y; x3[3 * i + 2] = p.z;
  if (type) type[i] = pos_type(p);
}
__global__ void k_update_pos(const double *__restrict__ x3, double4 *__restrict__ pos, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  p.x = x3[3 * i]; p.y = x3[3 * i + 1]; p.z = x3[3 * i + 2];
  pos[i] = p;
}

__global__ void k_soa6_to_aos(const double *__restrict__ src, size_t stride, double *__restrict__ dst, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
#pragma unroll
  for (int k = 0; k < 6; k++) dst[(size_t) i * 6 + k] = src[(size_t) k * stride + i];
}

static int stage_bytes(b200md_ctx *c, size_t bytes)
{
  return dev_ensure(c, c->stage, (bytes + 7) / 8);
}

extern "C" int b200md_upload_atoms(b200md_ctx *c, int n, const double *x, const double *v, const int *type, const int *tag,
                                   const int *image)
{
  NEED_CTX();
  if (n <= 0 || !x || !type || !tag) return b200md_fail(c, B200MD_ERR_ARG, "upload_atoms: n, x, type, tag are required");
  size_t est = (size_t) n + n / 4 + 1024;  // head-room for ghosts; grown on demand in borders()
  TRY(dev_ensure(c, c->pos, est));
  TRY(dev_ensure(c, c->tag, est));
  TRY(dev_ensure(c, c->image, (size_t) n));
  for (int k = 0; k < 3; k++) { TRY(dev_ensure(c, c->v[k], (size_t) n)); TRY(dev_ensure(c, c->f[k], (size_t) n)); }
  TRY(stage_bytes(c, (size_t) n * 28));
  double *sx = c->stage.p;
  int *st = (int *) (sx + (size_t) 3 * n);
  const int g = cdiv(n, 256);
  CU(cudaMemcpyAsync(sx, x, (size_t) n * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(st, type, (size_t) n * 4, cudaMemcpyHostToDevice, c->stream));
  k_pack_pos<<<g, 256, 0, c->stream>>>(sx, st, c->pos.p, n);
  LAUNCH_CHECK();
  if (v) {
    CU(cudaMemcpyAsync(sx, v, (size_t) n * 24, cudaMemcpyHostToDevice, c->stream));
    k_aos_to_soa<<<g, 256, 0, c->stream>>>(sx, c->v[0].p, c->v[1].p, c->v[2].p, n);
    LAUNCH_CHECK();
  } else
    for (int k = 0; k < 3; k++) CU(cudaMemsetAsync(c->v[k].p, 0, (size_t) n * 8, c->stream));
  for (int k = 0; k < 3; k++) CU(cudaMemsetAsync(c->f[k].p, 0, (size_t) n * 8, c->stream));
  CU(cudaMemcpyAsync(c->tag.p, tag, (size_t) n * 4, cudaMemcpyHostToDevice, c->stream));
  if (image) CU(cudaMemcpyAsync(c->image.p, image, (size_t) n * 4, cudaMemcpyHostToDevice, c->stream));
  else {
    // IMGMAX in each of the three fields == image (0,0,0) (src/atom_vec.cpp create_atom)
    std::vector<int> img((size_t) n, (IMGMAX << IMG2BITS) | (IMGMAX << IMGBITS) | IMGMAX);
    CU(cudaMemcpyAsync(c->image.p, img.data(), (size_t) n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  CU(cudaStreamSynchronize(c->stream));
  c->nlocal = n;
  c->nghost = 0;
  c->host_ghosts = false;  // the device makes the ghosts from here on (b200md_host_atoms switches back)
  c->list_valid = false;
  c->setup_done = false;
  c->have_ev = false;
  return B200MD_OK;
}

extern "C" int b200md_nlocal(const b200md_ctx *c) { return c ? c->nlocal : 0; }
extern "C" int b200md_nghost(const b200md_ctx *c) { return c ? c->nghost : 0; }

extern "C" int b200md_download_atoms(b200md_ctx *c, double *x, double *v, double *f, int *type, int *tag, int *image)
{
  NEED_CTX();
  const int n = c->nlocal;
  if (n <= 0) return b200md_fail(c, B200MD_ERR_ARG, "no atoms");
  TRY(stage_bytes(c, (size_t) n * 28));
  double *sx = c->stage.p;
  int *st = (int *) (sx + (size_t) 3 * n);
  const int g = cdiv(n, 256);
  if (x || type) {
    k_unpack_pos<<<g, 256, 0, c->stream>>>(c->pos.p, sx, st, n);
    LAUNCH_CHECK();
    if (x) CU(cudaMemcpyAsync(x, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    if (type) CU(cudaMemcpyAsync(type, st, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (v) {
    k_soa_to_aos<<<g, 256, 0, c->stream>>>(c->v[0].p, c->v[1].p, c->v[2].p, sx, n);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(v, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (f) {
    k_soa_to_aos<<<g, 256, 0, c->stream>>>(c->f[0].p, c->f[1].p, c->f[2].p, sx, n);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(f, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (tag) CU(cudaMemcpyAsync(tag, c->tag.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  if (image) CU(cudaMemcpyAsync(image, c->image.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

// ---- owned atoms back in the CALLER's order: the permutation by tag done on the device ----------------------------
__global__ void k_slot_fill(const int *__restrict__ host_tag, int n, int maxtag, int *__restrict__ slot, int *__restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int t = host_tag[i];
  if (t < 1 || t > maxtag) flags[3] = 1;
  else slot[t] = i;
}
// device atom s goes to row slot[tag[s]] of the output: WHAT 0 = position, 1 = three SoA arrays, 2 = one int array
template <int WHAT>
__global__ void k_by_slot(const int *__restrict__ tag, const int *__restrict__ slot, int maxtag, int n, const double4 *__restrict__ pos,
                          const double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ cc,
                          const int *__restrict__ ia, double *__restrict__ out3, int *__restrict__ outi, int *__restrict__ flags)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int t = tag[s];
  const int i = (t >= 1 && t <= maxtag) ? slot[t] : -1;
  if (i < 0) { flags[3] = 1; return; }   // a tag the caller does not hold
  if (WHAT == 0) {
    const double4 p = pos[s];
    out3[3 * (size_t) i] = p.x; out3[3 * (size_t) i + 1] = p.y; out3[3 * (size_t) i + 2] = p.z;
  } else if (WHAT == 1) {
    out3[3 * (size_t) i] = a[s]; out3[3 * (size_t) i + 1] = b[s]; out3[3 * (size_t) i + 2] = cc[s];
  } else {
    outi[i] = ia[s];
  }
}

extern "C" int b200md_download_atoms_by_tag(b200md_ctx *c, int n, const int *host_tag, int maxtag, double *x, double *v, double *f,
                                            int *image)
{
  NEED_CTX();
  if (n != c->nlocal || n <= 0) return b200md_fail(c, B200MD_ERR_ARG, "download_atoms_by_tag: the caller holds %d atoms, the device %d", n, c->nlocal);
  if (!host_tag || maxtag < 1) return b200md_fail(c, B200MD_ERR_ARG, "download_atoms_by_tag: host_tag[n] and maxtag are required");
  if ((int64_t) maxtag > 4 * (int64_t) n + 1024)
    return b200md_fail(c, B200MD_ERR_ARG, "download_atoms_by_tag: tags too sparse for a direct table (maxtag %d for %d atoms); use b200md_download_atoms", maxtag, n);
  // scratch: host tags [n] + slot table [maxtag + 1] (ints), then one 3 x n (or n) output at a time
  const size_t ints = (size_t) n + (size_t) maxtag + 1;
  const size_t head = (ints * 4 + 255) / 256 * 256;
  TRY(stage_bytes(c, head + (size_t) n * 24));
  int *d_htag = (int *) c->stage.p, *d_slot = d_htag + n;
  double *out3 = (double *) ((char *) c->stage.p + head);
  const int g = cdiv(n, 256);
  CU(cudaMemsetAsync(c->flags_d + 3, 0, sizeof(int), c->stream));
  CU(cudaMemcpyAsync(d_htag, host_tag, (size_t) n * 4, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemsetAsync(d_slot, 0xff, ((size_t) maxtag + 1) * 4, c->stream));
  k_slot_fill<<<g, 256, 0, c->stream>>>(d_htag, n, maxtag, d_slot, c->flags_d);
  LAUNCH_CHECK();
#define BY_SLOT(W, OUT3, OUTI, A, B, CC, IA) k_by_slot<W><<<g, 256, 0, c->stream>>>(c->tag.p, d_slot, maxtag, n, c->pos.p, A, B, CC, IA, OUT3, OUTI, c->flags_d)
  if (x) {
    BY_SLOT(0, out3, nullptr, nullptr, nullptr, nullptr, nullptr);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(x, out3, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (v) {
    BY_SLOT(1, out3, nullptr, c->v[0].p, c->v[1].p, c->v[2].p, nullptr);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(v, out3, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (f) {
    BY_SLOT(1, out3, nullptr, c->f[0].p, c->f[1].p, c->f[2].p, nullptr);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(f, out3, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (image) {
    BY_SLOT(2, nullptr, (int *) out3, nullptr, nullptr, nullptr, c->image.p);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(image, out3, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  }
#undef BY_SLOT
  CU(cudaMemcpyAsync(c->flags_h + 3, c->flags_d + 3, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (c->flags_h[3]) return b200md_fail(c, B200MD_ERR_ARG, "download_atoms_by_tag: the caller's tags and the device's do not match (a tag outside 1..maxtag or unknown to the caller)");
  return B200MD_OK;
}

extern "C" int b200md_download_all(b200md_ctx *c, double *x, int *tag)
{
  NEED_CTX();
  const int n = c->nlocal + c->nghost;
  if (n <= 0) return b200md_fail(c, B200MD_ERR_ARG, "no atoms");
  TRY(stage_bytes(c, (size_t) n * 24));
  if (x) {
    k_unpack_pos<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->stage.p, nullptr, n);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(x, c->stage.p, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
  }
  if (tag) CU(cudaMemcpyAsync(tag, c->tag.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

// ---- pair styles ---------------------------------------------------------------------------------
extern "C" int b200md_lj_init(b200md_ctx *c, int ntypes, const double *cutsq, const double *lj1, const double *lj2,
                              const double *lj3, const double *lj4, const double *offset, const double special_lj[4])
{
  NEED_CTX();
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES) return b200md_fail(c, B200MD_ERR_ARG, "ntypes %d out of range", ntypes);
  const int nt2 = (ntypes + 1) * (ntypes + 1);
  std::vector<double> h((size_t) 6 * nt2 + 4);   // + the four special_lj weights ([0] = 1: an ordinary neighbor)
  const double *src[6] = {cutsq, lj1, lj2, lj3, lj4, offset};
  double cutmax = 0.0;
  for (int t = 0; t < 6; t++)
    for (int k = 0; k < nt2; k++) h[(size_t) t * nt2 + k] = src[t][k];
  for (int i = 1; i <= ntypes; i++)
    for (int j = 1; j <= ntypes; j++) cutmax = fmax(cutmax, sqrt(cutsq[i * (ntypes + 1) + j]));
  h[(size_t) 6 * nt2] = 1.0;
  for (int k = 1; k < 4; k++) h[(size_t) 6 * nt2 + k] = special_lj ? special_lj[k] : 1.0;
  if (c->lj.d) CU(cudaFree(c->lj.d));
  CU(cudaMalloc((void **) &c->lj.d, h.size() * sizeof(double)));
  CU(cudaMemcpy(c->lj.d, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice));
  c->lj.ntypes = ntypes;
  for (int t = 0; t < 6; t++) c->lj.h_onetype[t] = src[t][(ntypes + 1) + 1];
  for (int k = 0; k < 4; k++) c->lj.special_lj[k] = special_lj ? special_lj[k] : 1.0;
  c->lj.special_lj[0] = 1.0;
  c->pair_style = 1;
  c->cutforce = cutmax;
  if (c->ntypes == 0) c->ntypes = ntypes;
  return B200MD_OK;
}

extern "C" int b200md_eam_init(b200md_ctx *c, int ntypes, int nrho, int nr, double rdrho, double rdr, double rhomax,
                               double cutforcesq, int nfrho, const double *frho_spline, int nrhor, const double *rhor_spline,
                               int nz2r, const double *z2r_spline, const int *type2frho, const int *type2rhor,
                               const int *type2z2r)
{
  NEED_CTX();
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES) return b200md_fail(c, B200MD_ERR_ARG, "ntypes %d out of range", ntypes);
  EAMTables &T = c->eam;
  if (T.frho) { cudaFree(T.frho); cudaFree(T.rhor); cudaFree(T.z2r); cudaFree(T.rhoA); cudaFree(T.tab1); cudaFree(T.type2frho); cudaFree(T.type2rhor); cudaFree(T.type2z2r); }
  T.ntypes = ntypes; T.nrho = nrho; T.nr = nr; T.nfrho = nfrho; T.nrhor = nrhor; T.nz2r = nz2r;
  T.rdrho = rdrho; T.rdr = rdr; T.rhomax = rhomax; T.cutforcesq = cutforcesq;
  const size_t nf = (size_t) nfrho * (nrho + 1) * 7, nrr = (size_t) nrhor * (nr + 1) * 7, nz = (size_t) nz2r * (nr + 1) * 7;
  const int nt1 = ntypes + 1;
  CU(cudaMalloc((void **) &T.frho, nf * 8));
  CU(cudaMalloc((void **) &T.rhor, nrr * 8));
  CU(cudaMalloc((void **) &T.z2r, nz * 8));
  CU(cudaMalloc((void **) &T.rhoA, (size_t) nrhor * (nr + 1) * sizeof(double4)));
  CU(cudaMalloc((void **) &T.type2frho, nt1 * 4));
  CU(cudaMalloc((void **) &T.type2rhor, nt1 * nt1 * 4));
  CU(cudaMalloc((void **) &T.type2z2r, nt1 * nt1 * 4));
  CU(cudaMemcpy(T.frho, frho_spline, nf * 8, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(T.rhor, rhor_spline, nrr * 8, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(T.z2r, z2r_spline, nz * 8, cudaMemcpyHostToDevice));
  std::vector<double> a((size_t) nrhor * (nr + 1) * 4);
  for (size_t k = 0; k < (size_t) nrhor * (nr + 1); k++)
    for (int q = 0; q < 4; q++) a[4 * k + q] = rhor_spline[7 * k + 3 + q];
  CU(cudaMemcpy(T.rhoA, a.data(), a.size() * 8, cudaMemcpyHostToDevice));
  {  // single-type fast path tables (first rhor / z2r table), SoA so that random knots spread over the smem banks
    const int n1 = nr + 1;
    std::vector<double> t1((size_t) 11 * n1);
    for (int m = 0; m < n1; m++) {
      for (int q = 0; q < 4; q++) t1[(size_t) q * n1 + m] = rhor_spline[7 * m + 3 + q];
      for (int q = 0; q < 3; q++) t1[(size_t) (4 + q) * n1 + m] = rhor_spline[7 * m + q];
      for (int q = 0; q < 4; q++) t1[(size_t) (7 + q) * n1 + m] = z2r_spline[7 * m + 3 + q];
    }
    CU(cudaMalloc((void **) &T.tab1, t1.size() * 8));
    CU(cudaMemcpy(T.tab1, t1.data(), t1.size() * 8, cudaMemcpyHostToDevice));
  }
  CU(cudaMemcpy(T.type2frho, type2frho, nt1 * 4, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(T.type2rhor, type2rhor, nt1 * nt1 * 4, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(T.type2z2r, type2z2r, nt1 * nt1 * 4, cudaMemcpyHostToDevice));
  c->pair_style = 2;
  c->cutforce = sqrt(cutforcesq);
  if (c->ntypes == 0) c->ntypes = ntypes;
  return B200MD_OK;
}

extern "C" int b200md_eam_download(b200md_ctx *c, double *rho, double *fp)
{
  NEED_CTX();
  if (c->pair_style != 2 || !c->rho.p) return b200md_fail(c, B200MD_ERR_ARG, "eam not active");
  if (rho) CU(cudaMemcpyAsync(rho, c->rho.p, (size_t) c->nlocal * 8, cudaMemcpyDeviceToHost, c->stream));
  if (fp) CU(cudaMemcpyAsync(fp, c->fp.p, (size_t) c->nlocal * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

// ---- neighbor ------------------------------------------------------------------------------------
extern "C" int b200md_neigh_config(b200md_ctx *c, int ntypes, const double *cutneighsq, double cutneighmax, double skin,
                                   int every, int delay, int check, double binsize)
{
  NEED_CTX();
  if (ntypes < 1 || ntypes > B200MD_MAXTYPES) return b200md_fail(c, B200MD_ERR_ARG, "ntypes %d out of range", ntypes);
  if (every < 1 || delay < 0) return b200md_fail(c, B200MD_ERR_ARG, "Illegal neigh_modify every/delay");
  const int nt2 = (ntypes + 1) * (ntypes + 1);
  for (int k = 0; k < nt2; k++) c->cutneighsq_h[k] = cutneighsq[k];
  TRY(dev_ensure(c, c->cutneighsq_d, (size_t) nt2));
  CU(cudaMemcpy(c->cutneighsq_d.p, c->cutneighsq_h, nt2 * 8, cudaMemcpyHostToDevice));
  c->ntypes = ntypes;
  c->cutneighmax = cutneighmax;
  c->skin = skin;
  c->triggersq = 0.25 * skin * skin;  // src/neighbor.cpp:344
  c->every = every; c->delay = delay; c->dist_check = check ? 1 : 0;
  c->binsize_user = binsize;
  c->have_neigh = true;
  c->list_valid = false;
  c->maxneigh = 0;
  return B200MD_OK;
}

static int neigh_decide(b200md_ctx *c, int *flag)
{
  // Neighbor::decide (src/neighbor.cpp:2293-2308)
  *flag = 0;
  c->ago++;
  if (c->ago >= c->delay && c->ago % c->every == 0) {
    if (!c->dist_check) { *flag = 1; return B200MD_OK; }
    int moved = 0;
    TRY(neigh_check_distance(c, &moved));
    if (moved && c->ago == (c->every > c->delay ? c->every : c->delay)) c->ndanger++;
    *flag = moved;
  }
  return B200MD_OK;
}

extern "C" int b200md_neigh_decide(b200md_ctx *c, int *flag)
{
  NEED_CTX();
  if (!c->list_valid) { *flag = 1; return B200MD_OK; }
  return neigh_decide(c, flag);
}

extern "C" int b200md_neigh_build(b200md_ctx *c)
{
  NEED_CTX();
  return neigh_rebuild(c);
}

// blocked-transposed device list -> the reference's per-atom rows (firstneigh[i][0..numneigh[i]), src/neigh_list.h:52-61)
__global__ void k_list_to_csr(const int *__restrict__ numneigh, const int *__restrict__ neigh, int maxneigh, int nlocal,
                              const long long *__restrict__ first, int *__restrict__ out)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int *nb = neigh + (size_t) (i >> 5) * TILE * maxneigh + (i & 31);
  int *o = out + first[i];
  const int n = numneigh[i];
  for (int k = 0; k < n; k++) o[k] = nb[(size_t) k * TILE] & CLJMASK;
}

extern "C" int b200md_neigh_export(b200md_ctx *c, int *numneigh, int *neighbors, int64_t *total)
{
  NEED_CTX();
  if (!c->list_valid) return b200md_fail(c, B200MD_ERR_ARG, "no neighbor list built");
  const int n = c->nlocal;
  std::vector<int> nn((size_t) n);
  CU(cudaMemcpyAsync(nn.data(), c->numneigh.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  std::vector<long long> first((size_t) n + 1);
  first[0] = 0;
  for (int i = 0; i < n; i++) first[i + 1] = first[i] + nn[i];
  if (numneigh) memcpy(numneigh, nn.data(), (size_t) n * 4);
  if (total) *total = first[n];
  if (neighbors && first[n] > 0) {
    long long *first_d = nullptr;
    int *out_d = nullptr;
    CU(cudaMalloc((void **) &first_d, ((size_t) n + 1) * 8));
    CU(cudaMalloc((void **) &out_d, (size_t) first[n] * 4));
    CU(cudaMemcpyAsync(first_d, first.data(), ((size_t) n + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    k_list_to_csr<<<cdiv(n, 128), 128, 0, c->stream>>>(c->numneigh.p, c->neigh.p, c->maxneigh, n, first_d, out_d);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(neighbors, out_d, (size_t) first[n] * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(first_d);
    cudaFree(out_d);
  }
  return B200MD_OK;
}

extern "C" int b200md_neigh_stats(b200md_ctx *c, int64_t *nbuilds, int64_t *ndanger, int *maxneigh)
{
  NEED_CTX();
  if (nbuilds) *nbuilds = c->nbuilds;
  if (ndanger) *ndanger = c->ndanger;
  if (maxneigh) *maxneigh = c->maxneigh_seen;
  return B200MD_OK;
}

extern "C" int b200md_halo_forward(b200md_ctx *c)
{
  NEED_CTX();
  if (!c->list_valid) return b200md_fail(c, B200MD_ERR_ARG, "halo_forward before the first neighbor build");
  return halo_forward_pos(c);
}

// ---- sections, force, integrate, thermo ------------------------------------------------------------
struct Sect {  // accumulates device time per Timer section when profiling is on
  b200md_ctx *c;
  int which;
  cudaEvent_t a = nullptr, b = nullptr;
  Sect(b200md_ctx *ctx, int w) : c(ctx), which(w)
  {
    if (c->profile >= 2) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, c->stream); }
  }
  ~Sect()
  {
    if (c->profile >= 2) {
      cudaEventRecord(b, c->stream);
      cudaEventSynchronize(b);
      float ms = 0;
      cudaEventElapsedTime(&ms, a, b);
      c->sect_ms[which] += ms;
      cudaEventDestroy(a);
      cudaEventDestroy(b);
    }
  }
};

// ---- overlap mode: halo on s_comm under the force kernels of the ghost-free tiles ----------------------------
// CommBrick::forward_comm + Pair::compute of a step that does not reneighbor and tallies nothing.  Order:
//   main  : ... initial_integrate | record ev_main
//   s_comm: wait ev_main, pack + NCCL send/recv of ghost positions, record ev_halo
//   main  : force of tiles without ghosts in their lists  | wait ev_halo | force of the other tiles
// EAM: the same around the rho pass, then the {x,y,z,fp} halo under the ghost-free tiles of the force pass.
static int overlap_streams(b200md_ctx *c)
{
  if (c->s_comm) return B200MD_OK;
  CU(cudaStreamCreateWithFlags(&c->s_comm, cudaStreamNonBlocking));
  CU(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming));
  return B200MD_OK;
}

static bool overlap_usable(const b200md_ctx *c)
{
  if (!c->overlap || !c->tiles_valid || !c->list_valid || c->host_ghosts) return false;
  if (c->pair_style == 2) return pair_eam_can_split(c);
  return c->pair_style == 1;
}

static int halo_on_comm_stream(b200md_ctx *c, double4 *arr)
{
  CU(cudaEventRecord(c->ev_main, c->stream));
  CU(cudaStreamWaitEvent(c->s_comm, c->ev_main, 0));
  c->hstream = c->s_comm;
  int st = halo_forward_vec4(c, arr);
  c->hstream = nullptr;
  TRY(st);
  CU(cudaEventRecord(c->ev_halo, c->s_comm));
  return B200MD_OK;
}

static int halo_and_force_overlapped(b200md_ctx *c)
{
  TRY(overlap_streams(c));
  TRY(halo_on_comm_stream(c, c->pos.p));
  if (c->pair_style == 1) {
    TRY(pair_lj_compute_part(c, 0));
    CU(cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
    TRY(pair_lj_compute_part(c, 1));
  } else {
    TRY(pair_eam_rho_part(c, 0));
    CU(cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
    TRY(pair_eam_rho_part(c, 1));
    TRY(halo_on_comm_stream(c, c->posfp.p));  // PairEAM's forward_comm of fp, as {x + shift, fp} records
    TRY(pair_eam_force_part(c, 0));
    CU(cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
    TRY(pair_eam_force_part(c, 1));
  }
  return B200MD_OK;
}

static int force_compute(b200md_ctx *c, int eflag, int vflag)
{
  if (!c->list_valid) return b200md_fail(c, B200MD_ERR_ARG, "pair_compute before the first neighbor build");
  if (c->pair_style == 1) TRY(pair_lj_compute(c, eflag, vflag));
  else if (c->pair_style == 2) TRY(pair_eam_compute(c, eflag, vflag));
  else return b200md_fail(c, B200MD_ERR_ARG, "no pair style defined");
  if (eflag || vflag) {
    CU(cudaMemcpyAsync(c->red_h, c->red_d, 9 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->eng_vdwl = c->red_h[0] + (c->pair_style == 2 ? c->red_h[8] : 0.0);
    for (int k = 0; k < 6; k++) c->virial[k] = c->red_h[1 + k];
    c->have_ev = true;
  }
  return B200MD_OK;
}

extern "C" int b200md_pair_compute(b200md_ctx *c, int eflag, int vflag, double *eng_vdwl, double *virial)
{
  NEED_CTX();
  {
    Sect s(c, SECT_PAIR);
    TRY(force_compute(c, eflag, vflag));
  }
  if ((eflag || vflag)) {
    if (eng_vdwl) *eng_vdwl = c->eng_vdwl;
    if (virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  }
  return B200MD_OK;
}

// eflag_atom / vflag_atom: a pass of its own after the force evaluation (pair_lj.cu, pair_eam.cu)
extern "C" int b200md_pair_peratom(b200md_ctx *c)
{
  NEED_CTX();
  if (!c->list_valid) return b200md_fail(c, B200MD_ERR_ARG, "pair_peratom before the first neighbor build");
  const int n = c->nlocal;
  const size_t stride = ((size_t) n + 31) / 32 * 32;
  TRY(dev_ensure(c, c->peratom, 7 * stride));
  c->peratom_stride = stride;
  c->peratom_n = 0;
  {
    Sect s(c, SECT_PAIR);
    if (c->pair_style == 1) TRY(pair_lj_peratom(c, c->peratom.p, stride));
    else if (c->pair_style == 2) TRY(pair_eam_peratom(c, c->peratom.p, stride));
    else return b200md_fail(c, B200MD_ERR_ARG, "no pair style defined");
  }
  c->peratom_n = n;
  return B200MD_OK;
}

extern "C" int b200md_peratom_download(b200md_ctx *c, double *eatom, double *vatom)
{
  NEED_CTX();
  const int n = c->peratom_n;
  if (n <= 0 || n != c->nlocal) return b200md_fail(c, B200MD_ERR_ARG, "peratom_download: call b200md_pair_peratom after the force evaluation first");
  if (eatom) CU(cudaMemcpyAsync(eatom, c->peratom.p, (size_t) n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (vatom) {
    // device [6][stride] -> host [n][6] through the staging buffer
    TRY(stage_bytes(c, (size_t) n * 48));
    k_soa6_to_aos<<<cdiv(n, 256), 256, 0, c->stream>>>(c->peratom.p + c->peratom_stride, c->peratom_stride, c->stage.p, n);
    LAUNCH_CHECK();
    CU(cudaMemcpyAsync(vatom, c->stage.p, (size_t) n * 48, cudaMemcpyDeviceToHost, c->stream));
  }
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

extern "C" int b200md_nve_initial(b200md_ctx *c, double dtv, double dtf)
{
  NEED_CTX();
  Sect s(c, SECT_MODIFY);
  return nve_initial(c, dtv, dtf);
}

extern "C" int b200md_nve_final(b200md_ctx *c, double dtf)
{
  NEED_CTX();
  Sect s(c, SECT_MODIFY);
  return nve_final(c, dtf);
}

extern "C" int b200md_thermo(b200md_ctx *c, double natoms_total, double *out)
{
  NEED_CTX();
  if (!c->have_ev) return b200md_fail(c, B200MD_ERR_ARG, "thermo needs a pair_compute with eflag/vflag set");
  double t = 0.0;
  TRY(thermo_ke(c, &t));
  double loc[8] = {c->eng_vdwl, t, c->virial[0], c->virial[1], c->virial[2], c->virial[3], c->virial[4], c->virial[5]};
  if (c->nranks > 1) {
    CU(cudaMemcpyAsync(c->red_d + 16, loc, sizeof(loc), cudaMemcpyHostToDevice, c->stream));
    TRY(comm_allreduce_sum(c, c->red_d + 16, 8));
    CU(cudaMemcpyAsync(loc, c->red_d + 16, sizeof(loc), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  // ComputePressure::virial_compute adds the bond virial to the pair virial (src/compute_pressure.cpp:314-341)
  double vbond = c->bonds.maxb ? c->bonds.virial_bond[0] + c->bonds.virial_bond[1] + c->bonds.virial_bond[2] : 0.0;
  if (c->bonds.maxb && c->nranks > 1) {
    CU(cudaMemcpyAsync(c->red_d + 16, &vbond, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    TRY(comm_allreduce_sum(c, c->red_d + 16, 1));
    CU(cudaMemcpyAsync(&vbond, c->red_d + 16, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  // ComputeTemp (src/compute_temp.cpp:55-100): dof = 3N - 3; ComputePressure (src/compute_pressure.cpp:228-260);
  // KE = 0.5*dof*boltz*T (src/thermo.cpp:1712-1717)
  const double dof = 3.0 * natoms_total - 3.0;
  const double tfactor = c->mvv2e / (dof * c->boltz);
  const double temp = loc[1] * tfactor;
  const double inv_volume = 1.0 / (c->prd[0] * c->prd[1] * c->prd[2]);
  out[0] = loc[0];
  out[1] = 0.5 * dof * c->boltz * temp;
  out[2] = temp;
  out[3] = (dof * c->boltz * temp + loc[2] + loc[3] + loc[4] + vbond) / 3.0 * inv_volume * c->nktv2p;
  for (int k = 0; k < 6; k++) out[4 + k] = loc[2 + k];
  return B200MD_OK;
}

// eng_vdwl / virial[6] of THIS rank from the most recent force evaluation that tallied them (the Pair members
// src/pair.h:41-44 that Compute pe / pressure read)
extern "C" int b200md_last_ev(b200md_ctx *c, double *eng_vdwl, double *virial)
{
  NEED_CTX();
  if (!c->have_ev) return b200md_fail(c, B200MD_ERR_ARG, "no force evaluation has tallied energy/virial yet");
  if (eng_vdwl) *eng_vdwl = c->eng_vdwl;
  if (virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  return B200MD_OK;
}

// ---- Verlet::setup / Verlet::run -----------------------------------------------------------------
extern "C" int b200md_setup(b200md_ctx *c)
{
  NEED_CTX();
  if (c->bonds.maxb && !c->bonds.fene_set) return b200md_fail(c, B200MD_ERR_ARG, "All bond coeffs are not set");
  // the bonded neighbor build implements `special_bonds fene` only: 1-2 partners excluded, 1-3 and 1-4 at full weight
  // (src/npair.h:125-158 with special_flag 0 1 1); any other weights given to b200md_lj_init would be silently ignored
  if (c->bonds.maxb && c->pair_style == 1 &&
      (c->lj.special_lj[1] != 0.0 || c->lj.special_lj[2] != 1.0 || c->lj.special_lj[3] != 1.0))
    return b200md_fail(c, B200MD_ERR_ARG, "bonded systems need special_bonds fene (special_lj = {1, 0, 1, 1}), got {%g, %g, %g, %g}",
                       c->lj.special_lj[0], c->lj.special_lj[1], c->lj.special_lj[2], c->lj.special_lj[3]);
  c->langevin.nextsort = c->ntimestep;  // Verlet::setup sorts unconditionally (src/verlet.cpp:121)
  {
    Sect s(c, SECT_NEIGH);
    TRY(neigh_rebuild(c));
  }
  c->nbuilds = 0;  // the build inside setup is not counted (Neighbor::init resets ncalls)
  c->ndanger = 0;
  {
    Sect s(c, SECT_PAIR);
    TRY(force_compute(c, 1, 1));
    TRY(bond_compute(c, 1));
  }
  if (c->langevin.on) {  // modify->setup -> FixLangevin::setup -> post_force (src/fix_langevin.cpp:359)
    Sect s(c, SECT_MODIFY);
    if (!c->langevin.bounds_set) c->langevin.beginstep = c->langevin.endstep = c->ntimestep;
    TRY(langevin_apply(c, c->ntimestep));
  }
  c->setup_done = true;
  return B200MD_OK;
}

extern "C" int b200md_run(b200md_ctx *c, int nsteps, double dt, int thermo_every)
{
  NEED_CTX();
  if (!c->setup_done) TRY(b200md_setup(c));
  const double dtv = dt, dtf = 0.5 * dt * c->ftm2v;  // FixNVE::init (src/fix_nve.cpp:51-58)
  struct FuseGuard {   // the fused-integration request never outlives this call, whatever path returns
    b200md_ctx *c;
    ~FuseGuard() { c->fuse.on = c->fuse.done = c->fuse.check = c->fuse.chain = c->check_ahead = false; }
  } fuse_guard{c};
  bool pending_final = false;                         // final_integrate of the previous step still to do
  bool ahead = false;   // ... or already done, with this step's initial_integrate, by the previous step's force kernel
  const bool extras = c->bonds.maxb > 0 || c->langevin.on;  // in.chain: bond forces / fix langevin after the pair force
  if (c->langevin.on) {
    if (!c->langevin.bounds_set) {  // Run::command (src/run.cpp:172-176): this call is the whole run
      c->langevin.beginstep = c->ntimestep;
      c->langevin.endstep = c->ntimestep + nsteps;
    }
    c->langevin.dt = dt;
  }
  for (int istep = 1; istep <= nsteps; istep++) {
    c->ntimestep++;
    // Integrate::ev_set (src/integrate.cpp:138-183): energy/virial on thermo steps and on the last step
    const int ev = (istep == nsteps) || (thermo_every > 0 && c->ntimestep % thermo_every == 0);
    {
      Sect s(c, SECT_MODIFY);
      if (ahead) {}
      else if (pending_final) TRY(nve_final_initial(c, dtv, dtf));
      else TRY(nve_initial(c, dtv, dtf));
      pending_final = ahead = false;
    }
    int nflag = 0;
    {
      Sect s(c, SECT_NEIGH);
      TRY(neigh_decide(c, &nflag));
    }
    // a step that neither tallies nor ends the run leaves its final_integrate to the next step's fused kernel -- or,
    // with fix nve alone, to the force kernel itself (INTEG in pair_lj.cu / pair_eam.cu), which writes the next
    // positions into pos_tmp (free between rebuilds; same capacity as pos)
    const bool fuse_step = c->fuse_nve && !ev && istep != nsteps;
    // (in.chain: the lj/cut kernel's epilogue also takes the bonds and the counter-mode thermostat; other pair styles
    // with bonds leave them to k_chain_post below)
    c->fuse.chain = extras && c->pair_style == 1 && chain_post_usable(c);
    c->fuse.on = fuse_step && (!extras || c->fuse.chain);
    c->fuse.done = false;
    c->fuse.dtv = dtv;
    c->fuse.dtf = dtf;
    {   // will the next step's Neighbor::decide look at the distances (ago is 0 after a rebuild on this step)?
      const int ago_next = (nflag ? 0 : c->ago) + 1;
      c->fuse.check = fuse_step && c->dist_check && ago_next >= c->delay && ago_next % c->every == 0;
      if (c->fuse.check) CU(cudaMemsetAsync(c->flags_d + 2, 0, sizeof(int), c->stream));
    }
    if (!nflag && !ev && overlap_usable(c)) {
      if (c->fuse.on) TRY(dev_ensure(c, c->pos_tmp, c->pos.n, false, false));
      Sect s(c, SECT_PAIR);
      TRY(halo_and_force_overlapped(c));
    } else {
      if (!nflag) {
        Sect s(c, SECT_COMM);
        TRY(halo_forward_pos(c));
      } else {
        Sect s(c, SECT_NEIGH);
        TRY(neigh_rebuild(c));
      }
      if (c->fuse.on) TRY(dev_ensure(c, c->pos_tmp, c->pos.n, false, false));
      Sect s(c, SECT_PAIR);
      TRY(force_compute(c, ev, ev));
    }
    c->fuse.on = false;
    if (c->fuse.done) {   // v and the next positions are written: the owned part of pos_tmp becomes pos (ghosts follow
      std::swap(c->pos, c->pos_tmp);   // by the next halo or rebuild)
      c->check_ahead = c->fuse.check;
      c->fuse.done = false;
      ahead = true;
      continue;
    }
    if (extras && fuse_step && chain_post_usable(c)) {   // bonds + thermostat + fix nve in one pass per atom (bond.cu)
      c->fuse.on = true;
      TRY(dev_ensure(c, c->pos_tmp, c->pos.n, false, false));
      TRY(chain_post_fused(c, c->ntimestep));
      c->fuse.on = false;
      std::swap(c->pos, c->pos_tmp);
      c->check_ahead = c->fuse.check;
      c->fuse.done = false;
      ahead = true;
      continue;
    }
    if (extras) {  // Verlet::run: force->bond->compute, then modify->post_force (src/verlet.cpp:325-347)
      TRY(bond_compute(c, ev));
      TRY(langevin_apply(c, c->ntimestep));
    }
    if (ev || istep == nsteps) {
      Sect s(c, SECT_MODIFY);
      TRY(nve_final(c, dtf));
    } else
      pending_final = true;
  }
  if (c->bonds.maxb)  // "Bad FENE bond" of the run's last steps (k_bond_fene raises flags_d[4]; BondFENE::compute :88 aborts)
    CU(cudaMemcpyAsync(c->flags_h + 4, c->flags_d + 4, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (c->bonds.maxb && c->flags_h[4]) return b200md_fail(c, B200MD_ERR_ARG, "Bad FENE bond");
  return B200MD_OK;
}

extern "C" int64_t b200md_timestep(const b200md_ctx *c) { return c ? c->ntimestep : 0; }
extern "C" int b200md_set_timestep(b200md_ctx *c, int64_t ntimestep)
{
  if (!c) return b200md_fail(nullptr, B200MD_ERR_ARG, "NULL context");
  if (ntimestep < 0) return b200md_fail(c, B200MD_ERR_ARG, "Timestep must be >= 0");  // reset_timestep (src/update.cpp:470)
  c->ntimestep = ntimestep;
  return B200MD_OK;
}

// ---- host-buffer mode ------------------------------------------------------------------------------
// One whole timestep with HOST x, v in and x, v, f out.  PCIe is the bound (120 B/atom against about 4.6 ms of
// kernels at 16.4 M atoms), so the step is pipelined over HS_CHUNKS atom ranges on three streams:
//   s_in : H2D of x, v of chunk c                              (copy engine, host -> device)
//   main : k_hs_in(c) = unpack + FixNVE::initial_integrate + pack of the new x; then decide / halo|rebuild / force;
//          then k_hs_out(c) = FixNVE::final_integrate + pack of v, f
//   s_out: D2H of the new x of chunk c (runs while later chunks still upload: PCIe is full duplex), later v, f
// On steps without an energy/virial tally the force kernel is launched per chunk as well, k_hs_out(c) right behind
// it, so v, f of chunk c go down while the forces of chunk c+1 are computed.
// A reneighboring step permutes the owned atoms (spatial sort), which invalidates the x that already went down in
// the old order: those steps send x again after the rebuild.
#define HS_CHUNKS 8
__global__ void k_hs_in(const double *__restrict__ x3, const double *__restrict__ v3, double4 *__restrict__ pos,
                        double *__restrict__ vx, double *__restrict__ vy, double *__restrict__ vz, const double *__restrict__ fx,
                        const double *__restrict__ fy, const double *__restrict__ fz, const double *__restrict__ mass,
                        double *__restrict__ xo3, int first, int n, double dtv, double dtf)
{
  int i = first + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= first + n) return;
  double4 p = pos[i];
  const double dtfm = dtf / mass[pos_type(p)];
  // same operation order as k_nve_initial
  const double a = __dadd_rn(v3[3 * i], __dmul_rn(dtfm, fx[i]));
  const double b = __dadd_rn(v3[3 * i + 1], __dmul_rn(dtfm, fy[i]));
  const double cz = __dadd_rn(v3[3 * i + 2], __dmul_rn(dtfm, fz[i]));
  vx[i] = a; vy[i] = b; vz[i] = cz;
  p.x = __dadd_rn(x3[3 * i], __dmul_rn(dtv, a));
  p.y = __dadd_rn(x3[3 * i + 1], __dmul_rn(dtv, b));
  p.z = __dadd_rn(x3[3 * i + 2], __dmul_rn(dtv, cz));
  pos[i] = p;
  xo3[3 * i] = p.x; xo3[3 * i + 1] = p.y; xo3[3 * i + 2] = p.z;
}

__global__ void k_hs_out(const double4 *__restrict__ pos, double *__restrict__ vx, double *__restrict__ vy,
                         double *__restrict__ vz, const double *__restrict__ fx, const double *__restrict__ fy,
                         const double *__restrict__ fz, const double *__restrict__ mass, double *__restrict__ vo3,
                         double *__restrict__ fo3, int first, int n, double dtf)
{
  int i = first + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= first + n) return;
  const double dtfm = dtf / mass[__double2loint(pos[i].w)];
  const double gx = fx[i], gy = fy[i], gz = fz[i];
  const double a = __dadd_rn(vx[i], __dmul_rn(dtfm, gx));
  const double b = __dadd_rn(vy[i], __dmul_rn(dtfm, gy));
  const double cz = __dadd_rn(vz[i], __dmul_rn(dtfm, gz));
  vx[i] = a; vy[i] = b; vz[i] = cz;
  vo3[3 * i] = a; vo3[3 * i + 1] = b; vo3[3 * i + 2] = cz;
  fo3[3 * i] = gx; fo3[3 * i + 1] = gy; fo3[3 * i + 2] = gz;
}

static int hs_streams(b200md_ctx *c)
{
  if (c->s_in) return B200MD_OK;
  CU(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
  for (int k = 0; k < 3 * HS_CHUNKS + 2; k++) CU(cudaEventCreateWithFlags(&c->hs_ev[k], cudaEventDisableTiming));
  return B200MD_OK;
}

// the plain form: copies and kernels in order on the one stream (default; verified on B200 at 1 and 2 ranks)
static int step_host_serial(b200md_ctx *c, double dt, int eflag, int vflag, int capacity, double *x, double *v, double *f,
                            int *tag, int *nlocal_out, double *eng_vdwl, double *virial)
{
  int n = c->nlocal;
  const double dtv = dt, dtf = 0.5 * dt * c->ftm2v;
  TRY(stage_bytes(c, (size_t) capacity * 48));
  double *sx = c->stage.p, *sv = sx + (size_t) 3 * capacity;
  int g = cdiv(n, 256);
  CU(cudaMemcpyAsync(sx, x, (size_t) n * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(sv, v, (size_t) n * 24, cudaMemcpyHostToDevice, c->stream));
  k_update_pos<<<g, 256, 0, c->stream>>>(sx, c->pos.p, n);
  LAUNCH_CHECK();
  k_aos_to_soa<<<g, 256, 0, c->stream>>>(sv, c->v[0].p, c->v[1].p, c->v[2].p, n);
  LAUNCH_CHECK();
  c->ntimestep++;
  TRY(nve_initial(c, dtv, dtf));
  int nflag = 0;
  TRY(neigh_decide(c, &nflag));
  if (!nflag) TRY(halo_forward_pos(c));
  else {
    TRY(neigh_rebuild(c));
    n = c->nlocal;  // several ranks: atoms migrated at the rebuild (Comm::exchange)
    if (n > capacity) return b200md_fail(c, B200MD_ERR_ARG, "step_host: %d atoms after migration exceed the caller's capacity %d", n, capacity);
    g = cdiv(n, 256);
  }
  TRY(force_compute(c, eflag, vflag));
  TRY(nve_final(c, dtf));
  k_unpack_pos<<<g, 256, 0, c->stream>>>(c->pos.p, sx, nullptr, n);
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(x, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
  k_soa_to_aos<<<g, 256, 0, c->stream>>>(c->v[0].p, c->v[1].p, c->v[2].p, sv, n);
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(v, sv, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));  // sx is reused for f below
  k_soa_to_aos<<<g, 256, 0, c->stream>>>(c->f[0].p, c->f[1].p, c->f[2].p, sx, n);
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(f, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
  if (tag) CU(cudaMemcpyAsync(tag, c->tag.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (nlocal_out) *nlocal_out = n;
  if (eflag || vflag) {
    if (eng_vdwl) *eng_vdwl = c->eng_vdwl;
    if (virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  }
  return B200MD_OK;
}

extern "C" int b200md_step_host(b200md_ctx *c, double dt, int eflag, int vflag, int capacity, double *x, double *v, double *f,
                                int *tag, int *nlocal_out, double *eng_vdwl, double *virial)
{
  NEED_CTX();
  if (!c->setup_done) return b200md_fail(c, B200MD_ERR_ARG, "step_host needs b200md_setup first");
  if (!x || !v || !f) return b200md_fail(c, B200MD_ERR_ARG, "step_host: x, v, f are required");
  if (c->bonds.maxb > 0 || c->langevin.on)
    return b200md_fail(c, B200MD_ERR_ARG, "step_host: bonded / fix langevin systems run through b200md_setup + b200md_run");
  int n = c->nlocal;
  if (capacity < n) return b200md_fail(c, B200MD_ERR_ARG, "step_host: capacity %d < nlocal %d", capacity, n);
  // default: the three-stream chunk pipeline below (B200, 16.4 M atoms: 527 against 405 Matom-steps/s serial)
  if (!c->host_pipeline) return step_host_serial(c, dt, eflag, vflag, capacity, x, v, f, tag, nlocal_out, eng_vdwl, virial);
  const double dtv = dt, dtf = 0.5 * dt * c->ftm2v;
  TRY(hs_streams(c));
  TRY(stage_bytes(c, (size_t) capacity * 120));
  const size_t cap3 = (size_t) 3 * capacity;
  double *sx = c->stage.p, *sv = sx + cap3, *xo = sv + cap3, *vo = xo + cap3, *fo = vo + cap3;
  const int nch = n >= c->hs_min_chunk ? HS_CHUNKS : 1;
  const int per = cdiv(cdiv(n, nch), 256) * 256;
  cudaEvent_t *e_in = c->hs_ev, *e_x = c->hs_ev + HS_CHUNKS, *e_o = c->hs_ev + 2 * HS_CHUNKS;
  cudaEvent_t e_start = c->hs_ev[3 * HS_CHUNKS], e_mid = c->hs_ev[3 * HS_CHUNKS + 1];
  // the copy streams start after whatever the main stream still has queued (staging buffers are reused)
  CU(cudaEventRecord(e_start, c->stream));
  CU(cudaStreamWaitEvent(c->s_in, e_start, 0));
  CU(cudaStreamWaitEvent(c->s_out, e_start, 0));
  {
    Sect s(c, SECT_MODIFY);
    for (int ch = 0; ch < nch; ch++) {
      const int first = ch * per, m = min(per, n - first);
      if (m <= 0) break;
      CU(cudaMemcpyAsync(sx + (size_t) 3 * first, x + (size_t) 3 * first, (size_t) m * 24, cudaMemcpyHostToDevice, c->s_in));
      CU(cudaMemcpyAsync(sv + (size_t) 3 * first, v + (size_t) 3 * first, (size_t) m * 24, cudaMemcpyHostToDevice, c->s_in));
      CU(cudaEventRecord(e_in[ch], c->s_in));
      CU(cudaStreamWaitEvent(c->stream, e_in[ch], 0));
      k_hs_in<<<cdiv(m, 256), 256, 0, c->stream>>>(sx, sv, c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->f[0].p, c->f[1].p,
                                                    c->f[2].p, c->mass_d.p, xo, first, m, dtv, dtf);
      LAUNCH_CHECK();
      CU(cudaEventRecord(e_x[ch], c->stream));
      CU(cudaStreamWaitEvent(c->s_out, e_x[ch], 0));
      CU(cudaMemcpyAsync(x + (size_t) 3 * first, xo + (size_t) 3 * first, (size_t) m * 24, cudaMemcpyDeviceToHost, c->s_out));
    }
  }
  c->ntimestep++;
  int nflag = 0;
  TRY(neigh_decide(c, &nflag));
  if (!nflag) TRY(halo_forward_pos(c));
  else {
    TRY(neigh_rebuild(c));
    // the owned atoms were wrapped, re-ordered and (several ranks) migrated: send x again in the new order, after
    // the stale copies
    n = c->nlocal;
    if (n > capacity) {
      cudaStreamSynchronize(c->s_out);
      return b200md_fail(c, B200MD_ERR_ARG, "step_host: %d atoms after migration exceed the caller's capacity %d", n, capacity);
    }
    k_unpack_pos<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, xo, nullptr, n);
    LAUNCH_CHECK();
    CU(cudaEventRecord(e_mid, c->stream));
    CU(cudaStreamWaitEvent(c->s_out, e_mid, 0));
    CU(cudaMemcpyAsync(x, xo, (size_t) n * 24, cudaMemcpyDeviceToHost, c->s_out));
  }
  // without a tally the force kernel runs range by range too, so that the download of one range's v, f is under the
  // force kernel of the next (EAM: passes A+B and the fp halo need all atoms; pass C is ranged)
  const bool ranged = !eflag && !vflag && nch > 1 && c->list_valid &&
                      (c->pair_style == 1 ? pair_lj_range_usable(c) : c->pair_style == 2 && pair_eam_range_usable(c));
  if (!ranged) TRY(force_compute(c, eflag, vflag));
  else if (c->pair_style == 2) TRY(pair_eam_before_ranges(c));
  const int per2 = cdiv(cdiv(n, nch), 512) * 512;  // n may have changed at the rebuild; a multiple of the force kernels' blocks
  {
    for (int ch = 0; ch < nch; ch++) {
      const int first = ch * per2, m = min(per2, n - first);
      if (m <= 0) break;
      if (ranged) {
        Sect s(c, SECT_PAIR);
        TRY(c->pair_style == 1 ? pair_lj_compute_range(c, first, m) : pair_eam_force_range(c, first, m));
      }
      Sect s(c, SECT_MODIFY);
      k_hs_out<<<cdiv(m, 256), 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->f[0].p, c->f[1].p, c->f[2].p,
                                                     c->mass_d.p, vo, fo, first, m, dtf);
      LAUNCH_CHECK();
      CU(cudaEventRecord(e_o[ch], c->stream));
      CU(cudaStreamWaitEvent(c->s_out, e_o[ch], 0));
      CU(cudaMemcpyAsync(v + (size_t) 3 * first, vo + (size_t) 3 * first, (size_t) m * 24, cudaMemcpyDeviceToHost, c->s_out));
      CU(cudaMemcpyAsync(f + (size_t) 3 * first, fo + (size_t) 3 * first, (size_t) m * 24, cudaMemcpyDeviceToHost, c->s_out));
    }
  }
  if (tag) CU(cudaMemcpyAsync(tag, c->tag.p, (size_t) n * 4, cudaMemcpyDeviceToHost, c->s_out));
  CU(cudaStreamSynchronize(c->s_out));
  CU(cudaStreamSynchronize(c->stream));
  if (nlocal_out) *nlocal_out = n;
  if (eflag || vflag) {
    if (eng_vdwl) *eng_vdwl = c->eng_vdwl;
    if (virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  }
  return B200MD_OK;
}

// ---- host-ghost mode: Pair::compute under the stock Verlet loop -------------------------------------------
// The caller (PairLJCutB200 / PairEAMB200 under `run_style verlet`) owns the atoms, the ghosts and the rebuild
// decision, exactly like the GPU package (src/GPU/pair_lj_cut_gpu.cpp:85-118): x of owned + ghost atoms comes up
// every call, f of the owned atoms goes down; the device list is rebuilt when the host says it reneighbored.
__global__ void k_aos_add(const double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ cc,
                          double *__restrict__ a3, int n)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  a3[3 * i] += a[i]; a3[3 * i + 1] += b[i]; a3[3 * i + 2] += cc[i];
}

extern "C" int b200md_host_atoms(b200md_ctx *c, int nlocal, int nghost, const double *x, const int *type,
                                 const double sublo[3], const double subhi[3], int rebuild)
{
  NEED_CTX();
  if (nlocal <= 0 || nghost < 0 || !x || !type) return b200md_fail(c, B200MD_ERR_ARG, "host_atoms: nlocal, x, type are required");
  if (!c->have_neigh) return b200md_fail(c, B200MD_ERR_ARG, "host_atoms before b200md_neigh_config");
  if (c->bonds.maxb > 0) return b200md_fail(c, B200MD_ERR_ARG, "host_atoms: bonded systems run through b200md_setup + b200md_run");
  const int nall = nlocal + nghost;
  if (nlocal != c->nlocal || nghost != c->nghost || !c->host_ghosts) rebuild = 1;
  TRY(dev_ensure(c, c->pos, (size_t) nall));
  for (int k = 0; k < 3; k++) TRY(dev_ensure(c, c->f[k], (size_t) nlocal));
  TRY(stage_bytes(c, (size_t) nall * 28));
  double *sx = c->stage.p;
  int *st = (int *) (sx + (size_t) 3 * nall);
  CU(cudaMemcpyAsync(sx, x, (size_t) nall * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(st, type, (size_t) nall * 4, cudaMemcpyHostToDevice, c->stream));
  k_pack_pos<<<cdiv(nall, 256), 256, 0, c->stream>>>(sx, st, c->pos.p, nall);
  LAUNCH_CHECK();
  c->nlocal = nlocal;
  c->nghost = nghost;
  c->host_ghosts = true;
  for (int k = 0; k < 3; k++) {
    if (!(subhi[k] > sublo[k])) return b200md_fail(c, B200MD_ERR_ARG, "host_atoms: subhi <= sublo in dim %d", k);
    c->sublo[k] = sublo[k];
    c->subhi[k] = subhi[k];
  }
  c->have_box = true;
  if (rebuild) {
    Sect s(c, SECT_NEIGH);
    TRY(neigh_setup_bins(c));
    CU(cudaMemsetAsync(c->flags_d, 0, 6 * sizeof(int), c->stream));
    TRY(neigh_build_lists(c));
  }
  return B200MD_OK;
}

// host-ghost mode, molecular systems: the special-bond lists of the owned atoms and the tags of owned + ghost atoms,
// as LAMMPS holds them (Atom::nspecial / special / tag, Neighbor::special_flag).  Call before the b200md_host_atoms that
// rebuilds the list; maxspecial == 0 turns the special handling off.  The list build then treats a neighbor whose tag
// is in the atom's lists exactly as NPair::find_special does (src/npair.h:125-158): dropped (flag 0), kept as an
// ordinary neighbor (flag 1) or kept with its class in bits 28-29 of the entry (flag 2), which the lj/cut kernel
// turns into the factor_lj of src/pair_lj_cut.cpp:103-105.
extern "C" int b200md_host_special(b200md_ctx *c, int nlocal, int nall, const int *tag, int maxspecial, const int *nspecial,
                                   const int *special, const int special_flag[4])
{
  NEED_CTX();
  SpecialState &S = c->special;
  if (maxspecial <= 0) {
    S.on = false;
    S.maxspecial = 0;
    return B200MD_OK;
  }
  if (nlocal <= 0 || nall < nlocal || !tag || !nspecial || !special || !special_flag)
    return b200md_fail(c, B200MD_ERR_ARG, "host_special: tag, nspecial, special and special_flag are required");
  if (c->bonds.maxb > 0) return b200md_fail(c, B200MD_ERR_ARG, "host_special: the context holds device-resident bonds (b200md_set_bonds)");
  for (int k = 1; k < 4; k++)
    if (special_flag[k] < 0 || special_flag[k] > 2) return b200md_fail(c, B200MD_ERR_ARG, "host_special: special_flag[%d] = %d", k, special_flag[k]);
  TRY(dev_ensure(c, c->tag, (size_t) nall));
  TRY(dev_ensure(c, S.n3, (size_t) nlocal * 3));
  TRY(dev_ensure(c, S.tags, (size_t) nlocal * maxspecial));
  CU(cudaMemcpyAsync(c->tag.p, tag, (size_t) nall * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(S.n3.p, nspecial, (size_t) nlocal * 3 * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(S.tags.p, special, (size_t) nlocal * maxspecial * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));   // the host arrays are LAMMPS's and may move
  S.maxspecial = maxspecial;
  S.flag[0] = 1;
  for (int k = 1; k < 4; k++) S.flag[k] = special_flag[k];
  S.on = true;
  c->list_valid = false;
  return B200MD_OK;
}

extern "C" int b200md_host_forces(b200md_ctx *c, double *f, int accumulate)
{
  NEED_CTX();
  const int n = c->nlocal;
  if (n <= 0 || !f) return b200md_fail(c, B200MD_ERR_ARG, "host_forces: no atoms / NULL f");
  TRY(stage_bytes(c, (size_t) n * 24));
  double *sx = c->stage.p;
  if (accumulate) {
    CU(cudaMemcpyAsync(sx, f, (size_t) n * 24, cudaMemcpyHostToDevice, c->stream));
    k_aos_add<<<cdiv(n, 256), 256, 0, c->stream>>>(c->f[0].p, c->f[1].p, c->f[2].p, sx, n);
  } else
    k_soa_to_aos<<<cdiv(n, 256), 256, 0, c->stream>>>(c->f[0].p, c->f[1].p, c->f[2].p, sx, n);
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(f, sx, (size_t) n * 24, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

extern "C" int b200md_host_eam_rho(b200md_ctx *c, int eflag, double *rho, double *fp, double *eng_embed)
{
  NEED_CTX();
  if (c->pair_style != 2) return b200md_fail(c, B200MD_ERR_ARG, "eam not active");
  if (!c->list_valid) return b200md_fail(c, B200MD_ERR_ARG, "host_eam_rho before the first neighbor build");
  Sect s(c, SECT_PAIR);
  TRY(pair_eam_rho(c, eflag ? 1 : 0));
  if (rho) CU(cudaMemcpyAsync(rho, c->rho.p, (size_t) c->nlocal * 8, cudaMemcpyDeviceToHost, c->stream));
  if (fp) CU(cudaMemcpyAsync(fp, c->fp.p, (size_t) c->nlocal * 8, cudaMemcpyDeviceToHost, c->stream));
  if (eflag) CU(cudaMemcpyAsync(c->red_h + 8, c->red_d + 8, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (eflag && eng_embed) *eng_embed = c->red_h[8];
  return B200MD_OK;
}

extern "C" int b200md_host_eam_force(b200md_ctx *c, const double *fp_all, int eflag, int vflag, double *eng_pair, double *virial)
{
  NEED_CTX();
  if (c->pair_style != 2) return b200md_fail(c, B200MD_ERR_ARG, "eam not active");
  if (!c->list_valid || !fp_all) return b200md_fail(c, B200MD_ERR_ARG, "host_eam_force: no list / NULL fp");
  Sect s(c, SECT_PAIR);
  const int ev = (eflag || vflag) ? 1 : 0;
  // only the ghosts' fp changed hands on the host (PairEAM::unpack_forward_comm, pair_eam.cpp:866-876)
  if (c->nghost > 0)
    CU(cudaMemcpyAsync(c->fp.p + c->nlocal, fp_all + c->nlocal, (size_t) c->nghost * 8, cudaMemcpyHostToDevice, c->stream));
  TRY(pair_eam_force(c, ev, true));
  if (ev) {
    CU(cudaMemcpyAsync(c->red_h, c->red_d, 7 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (eng_pair) *eng_pair = c->red_h[0];
    if (virial) for (int k = 0; k < 6; k++) virial[k] = c->red_h[1 + k];
  }
  return B200MD_OK;
}

// ---- instrumentation -------------------------------------------------------------------------------
extern "C" int b200md_timers(b200md_ctx *c, double ms[5], int64_t *kernel_launches, int reset)
{
  NEED_CTX();
  if (ms) for (int k = 0; k < 5; k++) ms[k] = c->sect_ms[k];
  if (kernel_launches) *kernel_launches = c->launches;
  if (reset) {
    for (int k = 0; k < 5; k++) c->sect_ms[k] = 0.0;
    c->launches = 0;
  }
  return B200MD_OK;
}

void kprof_drain(b200md_ctx *c)
{
  if (c->kev_used == 0) return;
  cudaStreamSynchronize(c->stream);
  for (size_t k = 0; k < c->kev_used; k++) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, c->kev[k].a, c->kev[k].b) == cudaSuccess) {
      c->kc_ms[c->kev[k].cls] += ms;
      c->kc_count[c->kev[k].cls]++;
    }
  }
  c->kev_used = 0;
}

extern "C" int b200md_kernel_times(b200md_ctx *c, double *ms, int64_t *count, int reset)
{
  NEED_CTX();
  kprof_drain(c);
  for (int k = 0; k < KC_N; k++) {
    if (ms) ms[k] = c->kc_ms[k];
    if (count) count[k] = c->kc_count[k];
    if (reset) { c->kc_ms[k] = 0.0; c->kc_count[k] = 0; }
  }
  return B200MD_OK;
}

extern "C" int b200md_profile(b200md_ctx *c, int level)
{
  NEED_CTX();
  kprof_drain(c);
  c->profile = level < 0 ? 0 : level;
  return B200MD_OK;
}

// execution knobs, per context; defaults are the measured winners.  Unknown key: B200MD_ERR_ARG.
//   "fuse_nve"      0/1  fix nve inside the force kernel on steps that neither tally nor end a run (default 1)
//   "overlap"       0/1  halo exchange on its own stream under the force kernels of the ghost-free tiles
//   "host_pipeline" 0/1  b200md_step_host as a three-stream chunk pipeline (0: serial copy - compute - copy)
//   "hs_min_chunk"  n    smallest system the pipeline splits into ranges (tests lower it)
//   "build_gb" n, "build_fp64" 0/1   test knobs of the list build: bins per block; force the all-FP64 fallback kernel
//   "build_bond_general" 0/1         test knob: bonded build with the closest-image choice even where images are unique
extern "C" int b200md_tune(b200md_ctx *c, const char *key, int value)
{
  NEED_CTX();
  if (!key) return b200md_fail(c, B200MD_ERR_ARG, "b200md_tune: key == NULL");
  if (!strcmp(key, "halo_put")) c->put.enabled = value != 0;   // takes effect with the next list build
  else if (!strcmp(key, "fuse_nve")) c->fuse_nve = value != 0;
  else if (!strcmp(key, "overlap")) c->overlap = value != 0;   // takes effect with the next list build (tile flags)
  else if (!strcmp(key, "host_pipeline")) c->host_pipeline = value != 0;
  else if (!strcmp(key, "hs_min_chunk")) c->hs_min_chunk = value > 0 ? value : 65536;
  else if (!strcmp(key, "build_gb")) c->build_gb_user = value > 0 ? value : 0;     // test knob
  else if (!strcmp(key, "build_fp64")) c->build_fp64 = value != 0;                 // test knob
  else if (!strcmp(key, "build_threads")) { c->build_threads = value; c->seg_cap = 0; }   // 128 or 256 threads per block of the list build
  else if (!strcmp(key, "build_bond_general")) c->build_bond_general = value != 0; // test knob
  else if (!strcmp(key, "build_seg_cap")) { c->seg_cap_user = value > 0 ? value : 0; c->seg_cap = 0; }   // test knob
  else return b200md_fail(c, B200MD_ERR_ARG, "b200md_tune: unknown key '%s'", key);
  return B200MD_OK;
}
extern "C" int b200md_build_bins_per_block(const b200md_ctx *c) { return c ? c->build_gb : 0; }
extern "C" void *b200md_stream(b200md_ctx *c) { return c ? (void *) c->stream : nullptr; }
