// nve.cu -- fix nve on the device: FixNVE::initial_integrate / final_integrate (src/fix_nve.cpp:64-141),
// and the kinetic-energy sum of ComputeTemp::compute_scalar (src/compute_temp.cpp:72-100).
// Streaming kernels; arithmetic is written with explicit round-to-nearest mul/add (no FMA contraction)
// so each atom's update is bit-identical to the reference's x86-64 code for the same inputs.
#include "ctx.cuh"

__global__ void k_nve_initial(double4 *__restrict__ pos, double *__restrict__ vx, double *__restrict__ vy,
                              double *__restrict__ vz, const double *__restrict__ fx, const double *__restrict__ fy,
                              const double *__restrict__ fz, const double *__restrict__ mass, int n, double dtv, double dtf)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  const double dtfm = dtf / mass[pos_type(p)];
  double a = __dadd_rn(vx[i], __dmul_rn(dtfm, fx[i]));
  double b = __dadd_rn(vy[i], __dmul_rn(dtfm, fy[i]));
  double cz = __dadd_rn(vz[i], __dmul_rn(dtfm, fz[i]));
  vx[i] = a;
  vy[i] = b;
  vz[i] = cz;
  p.x = __dadd_rn(p.x, __dmul_rn(dtv, a));
  p.y = __dadd_rn(p.y, __dmul_rn(dtv, b));
  p.z = __dadd_rn(p.z, __dmul_rn(dtv, cz));
  pos[i] = p;
}

__global__ void k_nve_final(const double4 *__restrict__ pos, double *__restrict__ vx, double *__restrict__ vy,
                            double *__restrict__ vz, const double *__restrict__ fx, const double *__restrict__ fy,
                            const double *__restrict__ fz, const double *__restrict__ mass, int n, double dtf)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double dtfm = dtf / mass[__double2loint(pos[i].w)];
  vx[i] = __dadd_rn(vx[i], __dmul_rn(dtfm, fx[i]));
  vy[i] = __dadd_rn(vy[i], __dmul_rn(dtfm, fy[i]));
  vz[i] = __dadd_rn(vz[i], __dmul_rn(dtfm, fz[i]));
}

// final_integrate of step n fused with initial_integrate of step n+1 (same per-atom operation order as
// the two separate kernels, one pass over v and f instead of two)
__global__ void k_nve_final_initial(double4 *__restrict__ pos, double *__restrict__ vx, double *__restrict__ vy,
                                    double *__restrict__ vz, const double *__restrict__ fx, const double *__restrict__ fy,
                                    const double *__restrict__ fz, const double *__restrict__ mass, int n, double dtv,
                                    double dtf)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos[i];
  const double dtfm = dtf / mass[pos_type(p)];
  const double gx = __dmul_rn(dtfm, fx[i]), gy = __dmul_rn(dtfm, fy[i]), gz = __dmul_rn(dtfm, fz[i]);
  double a = __dadd_rn(__dadd_rn(vx[i], gx), gx);
  double b = __dadd_rn(__dadd_rn(vy[i], gy), gy);
  double cz = __dadd_rn(__dadd_rn(vz[i], gz), gz);
  vx[i] = a;
  vy[i] = b;
  vz[i] = cz;
  p.x = __dadd_rn(p.x, __dmul_rn(dtv, a));
  p.y = __dadd_rn(p.y, __dmul_rn(dtv, b));
  p.z = __dadd_rn(p.z, __dmul_rn(dtv, cz));
  pos[i] = p;
}

__global__ void __launch_bounds__(256)
k_ke(const double4 *__restrict__ pos, const double *__restrict__ vx, const double *__restrict__ vy,
     const double *__restrict__ vz, const double *__restrict__ mass, int n, double *__restrict__ partials)
{
  __shared__ double s_red[8];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double t = 0.0;
  if (i < n) {
    double a = vx[i], b = vy[i], cz = vz[i];
    t = (a * a + b * b + cz * cz) * mass[__double2loint(pos[i].w)];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int k = 0; k < 8; k++) s += s_red[k];
    partials[(size_t) blockIdx.x * 8] = s;
  }
}

int nve_initial(b200md_ctx *c, double dtv, double dtf)
{
  const int n = c->nlocal;
  KProf kp(c, KC_NVE);
  k_nve_initial<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->f[0].p, c->f[1].p,
                                                      c->f[2].p, c->mass_d.p, n, dtv, dtf);
  LAUNCH_CHECK();
  return B200MD_OK;
}

int nve_final(b200md_ctx *c, double dtf)
{
  const int n = c->nlocal;
  KProf kp(c, KC_NVE);
  k_nve_final<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->f[0].p, c->f[1].p, c->f[2].p,
                                                    c->mass_d.p, n, dtf);
  LAUNCH_CHECK();
  return B200MD_OK;
}

int nve_final_initial(b200md_ctx *c, double dtv, double dtf)
{
  const int n = c->nlocal;
  KProf kp(c, KC_NVE);
  k_nve_final_initial<<<cdiv(n, 256), 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->f[0].p, c->f[1].p,
                                                            c->f[2].p, c->mass_d.p, n, dtv, dtf);
  LAUNCH_CHECK();
  return B200MD_OK;
}

// sum_i m_i v_i^2 over owned atoms -> red_d[9] -> host
int thermo_ke(b200md_ctx *c, double *sum_mv2)
{
  const int n = c->nlocal, blocks = cdiv(n, 256);
  TRY(dev_ensure(c, c->partials, (size_t) blocks * 8));
  k_ke<<<blocks, 256, 0, c->stream>>>(c->pos.p, c->v[0].p, c->v[1].p, c->v[2].p, c->mass_d.p, n, c->partials.p);
  LAUNCH_CHECK();
  TRY(reduce_partials(c, blocks, 1, 9));
  CU(cudaMemcpyAsync(c->red_h + 9, c->red_d + 9, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *sum_mv2 = c->red_h[9];
  return B200MD_OK;
}
