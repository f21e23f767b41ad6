// pcb200.cu -- C ABI (include/pcb200.h) on top of the kernels in kernels.cuh.
//
// Host-side orchestration of one RK substep mirrors pde()/time_step
// (equ.f90:24-668, timestep.f90:65-198) but keeps f resident on the device:
//   BC_x -> BC_y -> [z halo exchange over NCCL | BC_z] -> fused RHS+RK kernel
//   -> (substep 1 with ldt) max-reduction -> dt -> streaming update.
// No CPU fallback: every entry point fails loudly without a CUDA device.
#include "kernels.cuh"
#include "rhs_dly.cuh"
#include "particles.cuh"
#include "snapshot.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <dlfcn.h>
#include <string>
#include <vector>
#include <algorithm>

// ---- minimal NCCL declarations (resolved with dlopen so that single-GPU use
// ---- does not need libnccl; same ABI as nccl.h 2.x) -----------------------
extern "C" {
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
typedef int ncclDataType_t;   // ncclFloat64 = 8
typedef int ncclRedOp_t;      // ncclSum=0, ncclMax=2
}
struct NcclApi {
  void *h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t *, void *) = nullptr;     // NCCL >= 2.18; optional
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

struct Ctx {
  bool inited = false, have_par = false;
  int dev = 0;
  cudaStream_t st = nullptr, st_comm = nullptr;
  cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr, ev_k0 = nullptr, ev_k1 = nullptr, ev_bc = nullptr, ev_halo = nullptr, ev_inner = nullptr;
  pcb200_params P;
  DevPar hp;
  size_t nelem = 0;             // mw*mvar
  double *f[2] = {nullptr, nullptr}, *w = nullptr;
  CUtensorMap tmap[2];          // TMA view of f[0], f[1] as (mx, my, mz*mvar)
  CUtensorMap tmap4[2];         // rhs_dly: f[0], f[1] as (mx, my, mz, mvar), box {HX,HY,1,mvar}
  CUtensorMap tmap_wu, tmap_wr; // rhs_dly: w as (mx, my, mz, mvar), boxes {TX+2,TY,1,3} and {TX+2,TY,1,mvar-3} at x0+2 (16-byte aligned start)
  bool have_dly = false;        // standard variable order + 4-D maps encoded
  bool have_tma = false;
  int cur = 0;
  double *partials = nullptr; size_t npartials = 0;
  double *red_dev = nullptr;    // [NRED]
  double *dt_dev = nullptr;     // [0]=dt [1]=dt1
  double *pinned = nullptr;     // [NRED+2]
  double *zprof[PCB200_NPROF] = {nullptr, nullptr, nullptr};
  double *f_host = nullptr; bool f_host_pinned = false;
  bool have_forcing = false; pcb200_forcing frc;
  double dt = 0.0;
  double diag[PCB200_NDIAG];
  bool use_global = false;      // PCB200_KERNEL=global
  bool use_cpasync = false;     // PCB200_KERNEL=cpasync: cp.async staging instead of TMA
  bool use_dly = true;          // rhs_dly (delayed z-mixed terms, tensor-memory history); PCB200_KERNEL=tma selects rhs_tma
  int zchunk = 0;
  const char *kname = "";        // kernel of the last substep launch
  cudaStream_t bcs = nullptr;   // stream of the ghost-zone work (compute stream, or the comm stream while it overlaps interior tiles)
  int overlap_force = -1;       // PCB200_OVERLAP_XY=1: overlap the x/y face exchange with the interior tiles (off by default, see do_substep)
  bool zsplit = false;          // PCB200_ZSPLIT=1: interior planes overlap with the z halo exchange (3 launches per substep); default: wait for the halo, one launch -- faster since the kernels got shorter (28.4 vs 28.9 ms/step on 2 GPUs)
  bool use_prepass = true;      // isothermal runs: CFL pre-pass + fused update in substep 1 (PCB200_PREPASS=0: axpy pass)
  int zchunk_dly = 0;           // rhs_dly pays 6 extra steps per chunk: fewer, longer chunks
  bool me_tma = false;          // PCB200_ME_TMA=1: magnetic+entropy on rhs_tma (the kernel of round 1) instead of rhs_dly
  int64_t launches = 0;
  float rhs_ms = 0.f; int rhs_n = 0; bool time_rhs = false;
  // multi-GPU
  NcclApi nccl; ncclComm_t comm = nullptr; int rank = 0, nranks = 1;
  // One communicator per decomposed axis, holding the nproc[axis] ranks of this rank's line along that axis
  // (ncclCommSplit; rank in it = position along the axis).  Point-to-point traffic of an 8-rank communicator gets
  // a fraction of the channels per peer: measured at 2x2x2, 512^3 per GPU, 0.66 ms per x or y face exchange and
  // 1.4 ms for z, against 0.24-0.29 ms for the same messages in a 2-rank communicator.
  ncclComm_t comm_ax[3] = {nullptr, nullptr, nullptr}; int split_np[3] = {0, 0, 0}; bool use_split = true;
  // dust particles (particles.cuh)
  struct Part {
    bool on = false;
    pcb200_particle_params pp;
    PartPar q;
    int cur = 0, nimn = 0, key_bits = 1;
    double *fp[2] = {nullptr, nullptr}, *dfp[2] = {nullptr, nullptr};
    int *ipar[2] = {nullptr, nullptr}, *inear[2] = {nullptr, nullptr};
    unsigned *keys[2] = {nullptr, nullptr}, *vals[2] = {nullptr, nullptr};
    int *imn_array = nullptr, *k1 = nullptr, *k2 = nullptr;
    int ncell = 0, cell_bits = 1;
    unsigned *ckeys[2] = {nullptr, nullptr}, *cvals[2] = {nullptr, nullptr};   // second sort: by cell
    int *ccount = nullptr, *cs = nullptr;                                      // per-cell count, exclusive scan (ncell+1)
    double *rhop = nullptr, *drag = nullptr, *aux = nullptr, *dt1 = nullptr;
    // migration between ranks (migrate_particles)
    MigPar mg; int nbr[27]; int *slot_of_rank = nullptr, *counters = nullptr, *leave = nullptr, *pre = nullptr, *holes = nullptr, *tails = nullptr;
    double *sendbuf = nullptr, *recvbuf = nullptr; int *counters_host = nullptr;
    double *acc = nullptr;                       // multi-rank deposit: [4][mw] sums incl. the first ghost layer
    double *fold[4] = {nullptr, nullptr, nullptr, nullptr}; size_t fold_elems = 0;
    void *cub_tmp = nullptr; size_t cub_bytes = 0;
  } part;
  double *face[4] = {nullptr, nullptr, nullptr, nullptr};   // send lo/hi, recv lo/hi for the packed x,y faces
  size_t face_elems = 0;
  std::string err;
};
static Ctx g;
// PCB200_TIME_GHOST=1: CUDA events between the phases of the ghost-zone section of every substep (on the
// stream the phase runs on); pcb200_finalize prints the average per phase
static SnapWriter g_snap;
static bool g_time_ghost = false;
static std::vector<std::pair<int, cudaEvent_t>> g_ghost_ev;
enum { GM_START = 0, GM_PACK, GM_NCCL, GM_UNPACK, GM_Z, GM_BC, GM_END,
       // particle phases (same switch): around the gas kernel of a substep
       PM_FIRST, PM_MIG_K, PM_MIG_X, PM_MAP, PM_CELLS, PM_DRAG, PM_GAS, PM_DEPOSIT, PM_FOLD, PM_BACK, GM_N };
static const char *g_gm_name[GM_N] = {"start", "pack", "nccl x/y", "unpack", "z exchange", "bc/wrap", "wait",
                                      "p:first+wrap", "p:migrate kernels", "p:migrate exchange+sync", "p:map+sort+gather", "p:cell bins",
                                      "p:drag", "gas kernel etc", "p:deposit", "p:fold", "p:back-reaction"};
static void ghost_mark(int tag, cudaStream_t st)
{
  if (!g_time_ghost) return;
  cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); g_ghost_ev.push_back({tag, e});
}

static int fail(int code, const std::string &msg) { g.err = msg; return code; }
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return fail(PCB200_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
#define NK(call) do { ncclResult_t r_ = (call); if (r_ != 0) \
  return fail(PCB200_ENCCL, std::string(#call) + ": " + (g.nccl.GetErrorString ? g.nccl.GetErrorString(r_) : "nccl error")); } while (0)

extern "C" const char *pcb200_last_error(void) { return g.err.c_str(); }
extern "C" const char *pcb200_kernel_name(void) { return g.kname; }
extern "C" int pcb200_abi_version(void) { return PCB200_ABI_VERSION; }

extern "C" int pcb200_init(int local_rank)
{
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(PCB200_ENODEV, std::string("pcb200_init: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU fallback");
  g.dev = local_rank % ndev;
  CK(cudaSetDevice(g.dev));
  if (!g.inited) {
    CK(cudaStreamCreateWithFlags(&g.st, cudaStreamNonBlocking));
    {
      // comm stream at the highest priority: its small pack/unpack/NCCL kernels must get an SM as soon as one
      // frees up while the compute stream has thousands of one-CTA-per-SM blocks queued
      int plo = 0, phi = 0;
      CK(cudaDeviceGetStreamPriorityRange(&plo, &phi));
      CK(cudaStreamCreateWithPriority(&g.st_comm, cudaStreamNonBlocking, phi));
    }
    CK(cudaEventCreate(&g.ev_t0)); CK(cudaEventCreate(&g.ev_t1));
    CK(cudaEventCreate(&g.ev_k0)); CK(cudaEventCreate(&g.ev_k1));
    CK(cudaEventCreateWithFlags(&g.ev_bc, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_halo, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_inner, cudaEventDisableTiming));
    CK(cudaMalloc(&g.red_dev, sizeof(double) * NRED));
    CK(cudaMalloc(&g.dt_dev, sizeof(double) * 2));
    CK(cudaMallocHost(&g.pinned, sizeof(double) * (NRED + 2)));
    const char *k = getenv("PCB200_KERNEL");
    g.use_global = k && !strcmp(k, "global");
    g.use_cpasync = k && !strcmp(k, "cpasync");
    g.use_dly = !k || !strcmp(k, "dly");
    { const char *me = getenv("PCB200_ME_TMA"); g.me_tma = me && atoi(me) != 0; }
    { const char *sp = getenv("PCB200_COMM_SPLIT"); g.use_split = !sp || atoi(sp) != 0; }
    const char *pp = getenv("PCB200_PREPASS");
    g.use_prepass = !pp || atoi(pp) != 0;
    const char *zs = getenv("PCB200_ZSPLIT");
    g.zsplit = zs && atoi(zs) != 0;
    const char *ov = getenv("PCB200_OVERLAP_XY");
    g.overlap_force = ov ? (atoi(ov) != 0) : -1;
    g.bcs = g.st;
    { const char *tg = getenv("PCB200_TIME_GHOST"); g_time_ghost = tg && atoi(tg) != 0; }
    const char *t = getenv("PCB200_TIME_RHS");
    g.time_rhs = t && atoi(t) != 0;
    g.inited = true;
  }
  return PCB200_OK;
}

static void part_free();
static int split_axis_comms();
static void free_fields()
{
  g_snap.release();
  for (int i = 0; i < 2; i++) { if (g.f[i]) cudaFree(g.f[i]); g.f[i] = nullptr; }
  if (g.w) cudaFree(g.w); g.w = nullptr;
  if (g.partials) cudaFree(g.partials); g.partials = nullptr; g.npartials = 0;
}

// TMA descriptors: f seen as a 3-D tensor (mx, my, mz*mvar) of doubles, box {HX,HY,1}.
// rhs_dly: one CTA per SM, a chunk of zc planes takes zc+6 steps (+ ~6 steps of start-up latency);
// pick the chunk length that minimises waves x steps
static int dly_chunk(int tiles, int nz)
{
  int best = nz; double bestc = 1e30;
  for (int nch = 1; nch <= 64 && (nz + nch - 1) / nch >= 16; nch++) {
    const int z = (nz + nch - 1) / nch;
    const double waves = ceil((double)tiles * nch / 148.0);
    const double cost = waves * (z + 12);
    if (cost < bestc * 0.999) { bestc = cost; best = z; }
  }
  return best;
}

// cuTensorMapEncodeTiled is taken from the driver at run time (no -lcuda).
static int make_tensor_maps()
{
  g.have_tma = false;
  g.have_dly = false;
  const DevPar &h = g.hp;
  if ((h.mx * 8) % 16 != 0) return PCB200_OK;          // global strides must be multiples of 16 B: cp.async path instead
  typedef CUresult (*encode_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                               const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || !fn || qres != cudaDriverEntryPointSuccess) { cudaGetLastError(); return PCB200_OK; }
  encode_t encode = (encode_t)fn;
  for (int i = 0; i < 2; i++) {
    cuuint64_t dims[3] = {(cuuint64_t)h.mx, (cuuint64_t)h.my, (cuuint64_t)h.mz * h.mvar};
    cuuint64_t strides[2] = {(cuuint64_t)h.mx * 8, (cuuint64_t)h.sz * 8};
    cuuint32_t box[3] = {HX, HY, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(&g.tmap[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, g.f[i], dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(PCB200_ECUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  }
  g.have_tma = true;
  // rhs_dly: one 4-D box per plane, which needs the reference's standard registration order
  // (register.f90:124-135) and no further variables
  g.have_dly = false;
  const pcb200_params &P = g.P;
  const int nf = 4 + (P.lmagnetic ? 3 : 0) + (P.lentropy ? 1 : 0);
  bool std_order = P.iux == 1 && P.iuy == 2 && P.iuz == 3 && P.ilnrho == 4 && P.mvar == nf;
  if (P.lentropy) std_order = std_order && P.iss == 5;
  if (P.lmagnetic) { const int a0 = P.lentropy ? 6 : 5; std_order = std_order && P.iax == a0 && P.iay == a0 + 1 && P.iaz == a0 + 2; }
  if (!std_order) return PCB200_OK;
  {
    cuuint64_t dims[4] = {(cuuint64_t)h.mx, (cuuint64_t)h.my, (cuuint64_t)h.mz, (cuuint64_t)h.mvar};
    cuuint64_t strides[3] = {(cuuint64_t)h.mx * 8, (cuuint64_t)h.sz * 8, (cuuint64_t)h.mw * 8};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    for (int i = 0; i < 4; i++) {
      cuuint32_t box[4] = {HX, HY, 1, (cuuint32_t)h.mvar};
      CUtensorMap *tm = i < 2 ? &g.tmap4[i] : (i == 2 ? &g.tmap_wu : &g.tmap_wr);
      void *base = i < 2 ? (void *)g.f[i] : (void *)g.w;
      // w tiles: u (3 variables); the rest in one box, except magnetic+entropy: one-variable box (lnrho at plane n, ss at plane n-3)
      if (i >= 2) { box[0] = TX + 2; box[1] = TY; box[3] = i == 2 ? 3 : ((P.lmagnetic && P.lentropy) ? 1 : (cuuint32_t)(h.mvar - 3)); }
      CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) return fail(PCB200_ECUDA, "cuTensorMapEncodeTiled (4-D) failed (" + std::to_string((int)r) + ")");
    }
  }
  g.have_dly = true;
  return PCB200_OK;
}

extern "C" int pcb200_initialize(const pcb200_params *par)
{
  if (!g.inited) return fail(PCB200_ESTATE, "pcb200_initialize before pcb200_init");
  if (!par) return fail(PCB200_EARG, "null params");
  if (par->abi_version != PCB200_ABI_VERSION) return fail(PCB200_EARG, "ABI version mismatch");
  const pcb200_params &P = *par;
  // same policy as gpu_astaroth.f90:32-71: refuse what is not implemented
  if (!P.lhydro || !P.ldensity) return fail(PCB200_EUNSUPP, "No GPU implementation without hydro+density");
  if (!P.lbidiagonal_derij && ((P.nx + 2 * NG) % 2 != 0 || g.use_cpasync))
    return fail(PCB200_EUNSUPP, "lbidiagonal_derij=F needs the TMA kernel (even nx)");
  if (P.nx < 2 * NG || P.ny < 2 * NG || P.nz < 2 * NG) return fail(PCB200_EUNSUPP, "local grid smaller than 6 points per direction");
  if (P.mvar < 4 || P.mvar > PCB200_MAXVAR) return fail(PCB200_EARG, "mvar out of range");
  if (P.lentropy && !P.iss) return fail(PCB200_EARG, "lentropy without iss");
  if (P.lmagnetic && !P.iax) return fail(PCB200_EARG, "lmagnetic without iax");
  if (P.nprocx < 1 || P.nprocy < 1 || P.nprocz < 1 || P.nprocx * P.nprocy * P.nprocz != g.nranks)
    return fail(PCB200_EARG, "nprocx*nprocy*nprocz differs from the communicator size (call pcb200_comm_init first)");
  // rank layout of the reference: iproc = ipx + nprocx*(ipy + nprocy*ipz) (mpicomm.f90:367-369)
  if (P.ipx + P.nprocx * (P.ipy + P.nprocy * P.ipz) != g.rank || P.ipx < 0 || P.ipx >= P.nprocx || P.ipy < 0 || P.ipy >= P.nprocy)
    return fail(PCB200_EARG, "ipx,ipy,ipz do not match this rank (iproc = ipx + nprocx*(ipy + nprocy*ipz))");
  CK(cudaSetDevice(g.dev));
  part_free();          // particle buffers depend on the grid: pcb200_particles_initialize must follow
  const bool realloc = !g.have_par || g.P.nx != P.nx || g.P.ny != P.ny || g.P.nz != P.nz || g.P.mvar != P.mvar;
  g.P = P;
  { int rcs = split_axis_comms(); if (rcs) return rcs; }
  DevPar &h = g.hp;
  memset(&h, 0, sizeof(h));
  h.nx = P.nx; h.ny = P.ny; h.nz = P.nz; h.mx = P.nx + 2 * NG; h.my = P.ny + 2 * NG; h.mz = P.nz + 2 * NG;
  h.sy = h.mx; h.sz = (long long)h.mx * h.my; h.mw = h.sz * h.mz; h.mvar = P.mvar;
  for (int r = 0; r < NROLE; r++) h.slot[r] = -1;
  h.slot[R_UX] = P.iux - 1; h.slot[R_UY] = P.iuy - 1; h.slot[R_UZ] = P.iuz - 1; h.slot[R_LNRHO] = P.ilnrho - 1;
  if (P.lentropy) h.slot[R_SS] = P.iss - 1;
  if (P.lmagnetic) { h.slot[R_AX] = P.iax - 1; h.slot[R_AY] = P.iay - 1; h.slot[R_AZ] = P.iaz - 1; }
  for (int r = 0; r < NROLE; r++) if (h.slot[r] >= P.mvar) return fail(PCB200_EARG, "variable slot beyond mvar");
  for (int r = 0; r < NROLE; r++) { h.foff[r] = h.slot[r] >= 0 ? (long long)h.slot[r] * h.mw : 0; h.foff8[r] = h.foff[r] * 8; }
  h.lbidiag = P.lbidiagonal_derij;
  h.lupw_uu = P.lupw_uu; h.lupw_lnrho = P.lupw_lnrho; h.lupw_ss = P.lupw_ss;
  h.iheatcond = P.iheatcond; h.lcooling = P.lcooling; h.lviscosity_heat = P.lviscosity_heat; h.lohmic_heat = P.lohmic_heat;
  h.lgravz = P.lgravz; h.ldensity = P.ldensity; h.lhydro = P.lhydro;
  h.dx_1 = P.dx_1; h.dy_1 = P.dy_1; h.dz_1 = P.dz_1;
  h.dxyz_2 = h.dx_1 * h.dx_1 + h.dy_1 * h.dy_1 + h.dz_1 * h.dz_1;
  h.cdt = P.cdt; h.cdtv = P.cdtv; h.cdts = P.cdts;
  h.nu = P.nu; h.eta = P.eta; h.mu0 = P.mu0; h.mu01 = 1.0 / P.mu0;
  h.cs20 = P.cs20; h.gamma = P.gamma; h.gamma1 = 1.0 / P.gamma; h.gamma_m1 = P.gamma - 1.0;
  h.cp = P.cp; h.cp1 = 1.0 / P.cp; h.cv = h.gamma1 * P.cp; h.cv1 = P.gamma * h.cp1; h.lnrho0 = P.lnrho0;
  h.lnTT0 = h.gamma_m1 != 0.0 ? log(P.cs20 / (P.cp * h.gamma_m1)) : log(P.cs20 / P.cp);   // eos_idealgas.f90:230
  h.Omega2 = 2.0 * P.Omega; h.gravz = P.gravz; h.chi = P.chi; h.cool = P.cool; h.cs2cool = P.cs2cool;
  h.fx1 = h.dx_1 / 60.0; h.fy1 = h.dy_1 / 60.0; h.fz1 = h.dz_1 / 60.0;
  h.fx2 = h.dx_1 * h.dx_1; h.fy2 = h.dy_1 * h.dy_1; h.fz2 = h.dz_1 * h.dz_1;
  h.fxy = h.dx_1 * h.dy_1 / 720.0; h.fyz = h.dy_1 * h.dz_1 / 720.0; h.fzx = h.dz_1 * h.dx_1 / 720.0;
  {
    const double c1[3] = {45.0, -9.0, 1.0}, c2[4] = {-490.0 / 180.0, 270.0 / 180.0, -27.0 / 180.0, 2.0 / 180.0}, cm[3] = {270.0, -27.0, 2.0};
    for (int i = 0; i < 3; i++) { h.k1x[i] = h.fx1 * c1[i]; h.k1y[i] = h.fy1 * c1[i]; h.k1z[i] = h.fz1 * c1[i];
                                  h.kxy[i] = h.fxy * cm[i]; h.kyz[i] = h.fyz * cm[i]; h.kzx[i] = h.fzx * cm[i]; }
    for (int i = 0; i < 4; i++) { h.k2x[i] = h.fx2 * c2[i]; h.k2y[i] = h.fy2 * c2[i]; h.k2z[i] = h.fz2 * c2[i]; }
    for (int i = 0; i < 3; i++) {
      h.dlyu[i][0] = h.nu / 3.0 * cm[i] * h.fzx; h.dlyu[i][1] = h.nu / 3.0 * cm[i] * h.fyz;
      h.dlya[i][0] = h.mu01 * cm[i] * h.fzx;     h.dlya[i][1] = h.mu01 * cm[i] * h.fyz;
    }
    h.k2yz0 = h.k2y[0] + h.k2z[0]; h.k2zx0 = h.k2z[0] + h.k2x[0]; h.k2xy0 = h.k2x[0] + h.k2y[0]; h.k2xyz0 = h.k2x[0] + h.k2y[0] + h.k2z[0];
  }
  g.nelem = (size_t)h.mw * P.mvar;
  if (realloc) {
    free_fields();
    for (int i = 0; i < 2; i++) { CK(cudaMalloc(&g.f[i], g.nelem * sizeof(double))); CK(cudaMemsetAsync(g.f[i], 0, g.nelem * sizeof(double), g.st)); }
    CK(cudaMalloc(&g.w, g.nelem * sizeof(double)));
    CK(cudaMemsetAsync(g.w, 0, g.nelem * sizeof(double), g.st));
    for (int q = 0; q < PCB200_NPROF; q++) {
      if (g.zprof[q]) cudaFree(g.zprof[q]);
      CK(cudaMalloc(&g.zprof[q], sizeof(double) * P.nz));
      CK(cudaMemsetAsync(g.zprof[q], 0, sizeof(double) * P.nz, g.st));
    }
    g.cur = 0;
  }
  if (P.nprocx > 1 || P.nprocy > 1) {
    const size_t fx = P.nprocx > 1 ? (size_t)NG * h.my * h.mz * P.mvar : 0, fy = P.nprocy > 1 ? (size_t)NG * h.mx * h.mz * P.mvar : 0;
    const size_t need = fx > fy ? fx : fy;
    if (need > g.face_elems) {
      for (int i = 0; i < 4; i++) { if (g.face[i]) cudaFree(g.face[i]); CK(cudaMalloc(&g.face[i], need * sizeof(double))); }
      g.face_elems = need;
    }
  }
  for (int q = 0; q < PCB200_NPROF; q++) h.zprof[q] = g.zprof[q];
  {
    int rc = make_tensor_maps();
    if (rc) return rc;
  }
  // z chunking: enough CTAs to fill 148 SMs a few times over, chunks >= 32 planes
  {
    const int tiles = ((P.nx + TX - 1) / TX) * ((P.ny + TY - 1) / TY);
    int zc = P.nz;
    while (zc > 32 && tiles * ((P.nz + zc - 1) / zc) < 148 * 12) zc = (zc + 1) / 2;
    const char *e = getenv("PCB200_ZCHUNK");
    if (e && atoi(e) > 0) zc = atoi(e);
    g.zchunk = zc;
    g.zchunk_dly = dly_chunk(tiles, P.nz);
    if (e && atoi(e) > 0) g.zchunk_dly = atoi(e);
  }
  {
    const int zcmin = g.zchunk < g.zchunk_dly ? g.zchunk : g.zchunk_dly;
    const size_t nb_m = (size_t)((P.nx + TX - 1) / TX) * ((P.ny + TY - 1) / TY) * ((P.nz + zcmin - 1) / zcmin + 2);
    const size_t nb_g = (size_t)((P.nx + 31) / 32) * ((P.ny + 3) / 4) * ((P.nz + 1) / 2 + 2);
    const size_t nb_p = (size_t)((P.nx + TX - 1) / TX) * ((P.ny + TY - 1) / TY) * ((P.nz + 15) / 16);   // cfl_prepass
    const size_t nb_s = (size_t)((P.nx + TX - 1) / TX) * ((P.ny + TY - 1) / TY) * 70;                  // split launches, up to 64 chunks each
    const size_t need = std::max(std::max(std::max(nb_m, nb_g), nb_p), nb_s) * NRED;
    if (need > g.npartials) {
      if (g.partials) cudaFree(g.partials);
      CK(cudaMalloc(&g.partials, need * sizeof(double))); g.npartials = need;
      CK(cudaMemsetAsync(g.partials, 0, need * sizeof(double), g.st));     // dt-only steps write slot RD_DT1 only; reduce_partials reads all slots
    }
  }
  CK(cudaMemcpyToSymbolAsync(c_par, &h, sizeof(DevPar), 0, cudaMemcpyHostToDevice, g.st));
  // opt in to the large dynamic shared memory of the marching kernels
#define OPTIN(M, E, A) CK(cudaFuncSetAttribute(rhs_march<M, E, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MarchCfg<M, E>::SMEM))
  OPTIN(false, false, 0); OPTIN(false, false, 1); OPTIN(false, false, 2);
  OPTIN(true, false, 0);  OPTIN(true, false, 1);  OPTIN(true, false, 2);
  OPTIN(false, true, 0);  OPTIN(false, true, 1);  OPTIN(false, true, 2);
  OPTIN(true, true, 0);   OPTIN(true, true, 1);   OPTIN(true, true, 2);
#undef OPTIN
#define OPTINT(M, E, A, R) CK(cudaFuncSetAttribute(rhs_tma<M, E, A, R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TmaCfg<M, E>::SMEM)); \
                           CK(cudaFuncSetAttribute(rhs_tma<M, E, A, R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TmaCfg<M, E>::SMEM))
#define OPTINT4(M, E) OPTINT(M, E, 0, false); OPTINT(M, E, 1, false); OPTINT(M, E, 2, false); OPTINT(M, E, 0, true)
  OPTINT4(false, false); OPTINT4(true, false); OPTINT4(false, true); OPTINT4(true, true);
#undef OPTINT4
#undef OPTINT
#define OPTIND1(M, E, U) CK(cudaFuncSetAttribute(rhs_dly<M, E, 0, false, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DlyCfg<M, E>::SMEM)); \
                         CK(cudaFuncSetAttribute(rhs_dly<M, E, 1, false, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DlyCfg<M, E>::SMEM)); \
                         CK(cudaFuncSetAttribute(rhs_dly<M, E, 0, true, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DlyCfg<M, E>::SMEM))
#define OPTIND(M, E) OPTIND1(M, E, false); OPTIND1(M, E, true)
  OPTIND(false, false); OPTIND(true, false); OPTIND(false, true); OPTIND(true, true);
#undef OPTIND1
#undef OPTIND
  CK(cudaStreamSynchronize(g.st));
  g.have_par = true;
  return PCB200_OK;
}

extern "C" int pcb200_set_zprofile(int which, const double *prof, int nz)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_set_zprofile before pcb200_initialize");
  if (which < 0 || which >= PCB200_NPROF || !prof || nz != g.P.nz) return fail(PCB200_EARG, "bad z profile");
  CK(cudaMemcpyAsync(g.zprof[which], prof, sizeof(double) * nz, cudaMemcpyHostToDevice, g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

extern "C" int pcb200_register(double *f_host)
{
  if (!f_host) return fail(PCB200_EARG, "null f");
  if (g.f_host_pinned && g.f_host != f_host) { cudaHostUnregister(g.f_host); g.f_host_pinned = false; }
  g.f_host = f_host;      // aliased only (astaroth/gpu_astaroth.cc:224-234)
  return PCB200_OK;
}

// page-lock the registered host array once so that uploads/downloads run at
// full PCIe rate (the reference notes this as time-critical, astaroth/loadStore.cc:82)
static void pin_registered()
{
  if (g.f_host && !g.f_host_pinned && g.nelem) {
    const char *e = getenv("PCB200_PIN_HOST");
    if (e && !atoi(e)) return;
    if (cudaHostRegister(g.f_host, g.nelem * sizeof(double), cudaHostRegisterDefault) == cudaSuccess) g.f_host_pinned = true;
    else cudaGetLastError();   // not fatal: fall back to pageable copies
  }
}

extern "C" int pcb200_load_farray(const double *f_host)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_load_farray before pcb200_initialize");
  const double *src = f_host ? f_host : g.f_host;
  if (!src) return fail(PCB200_ESTATE, "no host array registered");
  if (src == g.f_host) pin_registered();
  CK(cudaMemcpyAsync(g.f[g.cur], src, g.nelem * sizeof(double), cudaMemcpyHostToDevice, g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

extern "C" int pcb200_set_forcing(const pcb200_forcing *frc)
{
  g.have_forcing = frc != nullptr;
  if (frc) g.frc = *frc;
  return PCB200_OK;
}

// ------------------------------------------------------------------------
// kernel-time bookkeeping: event pairs around every fused RHS launch,
// resolved lazily in pcb200_rhs_kernel_time (no sync in the hot loop)
// ------------------------------------------------------------------------
static std::vector<cudaEvent_t> g_evpool;
static size_t g_evused = 0;
static bool g_time_on = false;
extern "C" int pcb200_enable_kernel_timing(int on) { g_time_on = on != 0; g_evused = 0; return PCB200_OK; }
static cudaEvent_t next_event()
{
  if (g_evused == g_evpool.size()) { cudaEvent_t e; cudaEventCreate(&e); g_evpool.push_back(e); }
  return g_evpool[g_evused++];
}
extern "C" int pcb200_rhs_kernel_time(float *avg_ms, int *count)
{
  CK(cudaStreamSynchronize(g.st));
  float tot = 0.f; int n = 0;
  for (size_t i = 0; i + 1 < g_evused; i += 2) { float ms = 0.f; CK(cudaEventElapsedTime(&ms, g_evpool[i], g_evpool[i + 1])); tot += ms; n++; }
  g_evused = 0;
  if (avg_ms) *avg_ms = n ? tot / n : 0.f;
  if (count) *count = n;
  return PCB200_OK;
}

// ------------------------------------------------------------------------
static int launch_bc(double *f, int axis, int do_lo, int do_hi)
{
  const pcb200_params &P = g.P; const DevPar &h = g.hp;
  BcArgs b; memset(&b, 0, sizeof(b));
  b.f = f; b.axis = axis; b.do_face[0] = do_lo; b.do_face[1] = do_hi;
  const int32_t (*src)[PCB200_MAXVAR] = axis == 0 ? P.bcx : (axis == 1 ? P.bcy : P.bcz);
  for (int fc = 0; fc < 2; fc++) for (int k = 0; k < PCB200_MAXVAR; k++) b.code[fc][k] = src[fc][k];
  b.ilnrho = P.ilnrho - 1; b.iss = P.iss - 1;
  b.cp = h.cp; b.cv = h.cv; b.cv1 = h.cv1; b.gamma_m1 = h.gamma_m1; b.cs20 = h.cs20; b.lnrho0 = h.lnrho0;
  b.cs2top = P.cs2top; b.Fbot = P.Fbot; b.chi = P.chi;
  b.FbotKbot = (P.hcond0 * P.hcond1 != 0.0) ? P.Fbot / (P.hcond0 * P.hcond1) : 0.0;    // entropy.f90:721-725
  b.chiconst = P.iheatcond == PCB200_HCOND_CHICONST;
  for (int i = 0; i < 7; i++) b.dz2_bound[i] = P.dz2_bound[i];
  const int mm[3] = {h.mx, h.my, h.mz};
  const int nb = axis == 0 ? mm[1] : mm[0], nc = axis == 2 ? mm[1] : mm[2];
  dim3 blk(128), grd((nb + 127) / 128, nc);
  bc_fill<<<grd, blk, 0, g.bcs>>>(b);
  g.launches++;
  CK(cudaGetLastError());
  return PCB200_OK;
}

// periodic directions that wrap inside this rank, all in one launch
static bool wraps_locally(int axis)
{
  const pcb200_params &P = g.P;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz};
  if (!P.lperi[axis] || np[axis] != 1) return false;
  const int32_t (*src)[PCB200_MAXVAR] = axis == 0 ? P.bcx : (axis == 1 ? P.bcy : P.bcz);
  for (int fc = 0; fc < 2; fc++) for (int k = 0; k < P.mvar; k++) if (src[fc][k] != PCB200_BC_P) return false;
  return true;
}
static int launch_bc_periodic(double *f, bool wx, bool wy, bool wz)
{
  const pcb200_params &P = g.P; const DevPar &h = g.hp;
  BcPerArgs b; b.f = f; b.wrap[0] = wx; b.wrap[1] = wy; b.wrap[2] = wz;
  b.n0 = wx ? 6LL * h.my * h.mz : 0; b.n1 = wy ? 6LL * h.mx * h.mz : 0; b.n2 = wz ? 6LL * h.mx * h.my : 0;
  const long long n = b.n0 + b.n1 + b.n2;
  dim3 blk(256), grd((unsigned)((n + 255) / 256), P.mvar);
  bc_periodic<<<grd, blk, 0, g.bcs>>>(b);
  g.launches++;
  CK(cudaGetLastError());
  return PCB200_OK;
}

// neighbour ranks along one axis; -1 = physical boundary (mpicomm.f90:367-420)
static void neighbours(int axis, int *nlo, int *nhi, bool *first, bool *last)
{
  const pcb200_params &P = g.P;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz};
  int ip[3] = {P.ipx, P.ipy, P.ipz};
  const bool per = P.lperi[axis] != 0;
  const int me = ip[axis];
  auto rank_of = [&](int i) { int q[3] = {ip[0], ip[1], ip[2]}; q[axis] = i; return q[0] + np[0] * (q[1] + np[1] * q[2]); };
  *first = me == 0; *last = me == np[axis] - 1;
  *nlo = me > 0 ? rank_of(me - 1) : (per ? rank_of(np[axis] - 1) : -1);
  *nhi = me < np[axis] - 1 ? rank_of(me + 1) : (per ? rank_of(0) : -1);
}

// communicator and peer ranks of the halo traffic along one axis: the axis communicator (ranks = positions along
// the axis) when it exists, else the parent communicator with the global ranks of neighbours()
static ncclComm_t axis_comm(int axis, int *plo, int *phi)
{
  const pcb200_params &P = g.P;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz}, ip[3] = {P.ipx, P.ipy, P.ipz};
  int nlo, nhi; bool first, last;
  neighbours(axis, &nlo, &nhi, &first, &last);
  if (!g.comm_ax[axis]) { *plo = nlo; *phi = nhi; return g.comm; }
  *plo = nlo < 0 ? -1 : (ip[axis] + np[axis] - 1) % np[axis];
  *phi = nhi < 0 ? -1 : (ip[axis] + 1) % np[axis];
  return g.comm_ax[axis];
}

// (re)build the axis communicators for the current process layout; collective over all ranks
static int split_axis_comms()
{
  const pcb200_params &P = g.P;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz}, ip[3] = {P.ipx, P.ipy, P.ipz};
  if (g.nranks == 1 || !g.comm) return PCB200_OK;
  if (g.split_np[0] == np[0] && g.split_np[1] == np[1] && g.split_np[2] == np[2]) return PCB200_OK;
  for (int a = 0; a < 3; a++) { if (g.comm_ax[a] && g.nccl.CommDestroy) g.nccl.CommDestroy(g.comm_ax[a]); g.comm_ax[a] = nullptr; }
  g.split_np[0] = np[0]; g.split_np[1] = np[1]; g.split_np[2] = np[2];
  if (!g.use_split || !g.nccl.CommSplit) return PCB200_OK;
  for (int a = 0; a < 3; a++) {
    if (np[a] == 1 || np[a] == g.nranks) continue;       // not decomposed / the line is the whole communicator
    const int b = (a + 1) % 3, c = (a + 2) % 3;
    const int color = ip[b] + np[b] * ip[c];               // one colour per line along axis a
    NK(g.nccl.CommSplit(g.comm, color, ip[a], &g.comm_ax[a], nullptr));
  }
  return PCB200_OK;
}

// z halo exchange between slabs (initiate/finalize_isendrcv_bdry, mpicomm.f90:1325-2047,
// z direction; x,y ghosts ride along because whole mx*my planes move: zero packing)
static int exchange_z(double *f)
{
  const pcb200_params &P = g.P; const DevPar &h = g.hp;
  int nlo, nhi;
  ncclComm_t cm = axis_comm(2, &nlo, &nhi);
  const size_t cnt = (size_t)NG * h.sz;
  NK(g.nccl.GroupStart());
  for (int k = 0; k < P.mvar; k++) {
    double *b = f + (size_t)k * h.mw;
    if (nlo >= 0) NK(g.nccl.Send(b + (size_t)NG * h.sz, cnt, 8, nlo, cm, g.st_comm));            // my lowest interior planes
    if (nhi >= 0) NK(g.nccl.Recv(b + (size_t)(NG + P.nz) * h.sz, cnt, 8, nhi, cm, g.st_comm));   // upper ghosts
    if (nhi >= 0) NK(g.nccl.Send(b + (size_t)P.nz * h.sz, cnt, 8, nhi, cm, g.st_comm));          // my highest interior planes
    if (nlo >= 0) NK(g.nccl.Recv(b, cnt, 8, nlo, cm, g.st_comm));                                // lower ghosts
  }
  NK(g.nccl.GroupEnd());
  return PCB200_OK;
}

// x or y halo exchange: pack NG planes (full extent of the other axes) -> send/recv -> unpack,
// all on the compute stream (it must finish before any stencil runs; only z overlaps)
static int exchange_packed(double *f, int axis)
{
  const pcb200_params &P = g.P; const DevPar &h = g.hp;
  int nlo, nhi;
  ncclComm_t cm = axis_comm(axis, &nlo, &nhi);
  const int n = axis == 0 ? P.nx : P.ny;
  const long long per_var = axis == 0 ? (long long)NG * h.my * h.mz : (long long)NG * h.mx * h.mz;
  const size_t cnt = (size_t)per_var * P.mvar;
  // both faces per launch: pack (lower/upper interior planes) -> one NCCL group -> unpack (upper/lower ghosts)
  dim3 blk(256), grd((unsigned)((2 * per_var + 255) / 256), P.mvar);
  Face2Args a; a.f = f; a.axis = axis; a.n = n; a.do_lo = nlo >= 0; a.do_hi = nhi >= 0;
  ghost_mark(GM_BC, g.bcs);
  a.buf_lo = g.face[0]; a.buf_hi = g.face[1]; a.unpack = 0;
  face_copy2<<<grd, blk, 0, g.bcs>>>(a);
  g.launches++;
  CK(cudaGetLastError());
  ghost_mark(GM_PACK, g.bcs);
  NK(g.nccl.GroupStart());
  if (nlo >= 0) NK(g.nccl.Send(g.face[0], cnt, 8, nlo, cm, g.bcs));
  if (nhi >= 0) NK(g.nccl.Recv(g.face[3], cnt, 8, nhi, cm, g.bcs));
  if (nhi >= 0) NK(g.nccl.Send(g.face[1], cnt, 8, nhi, cm, g.bcs));
  if (nlo >= 0) NK(g.nccl.Recv(g.face[2], cnt, 8, nlo, cm, g.bcs));
  NK(g.nccl.GroupEnd());
  ghost_mark(GM_NCCL, g.bcs);
  a.buf_lo = g.face[2]; a.buf_hi = g.face[3]; a.unpack = 1;
  face_copy2<<<grd, blk, 0, g.bcs>>>(a);
  g.launches++;
  CK(cudaGetLastError());
  ghost_mark(GM_UNPACK, g.bcs);
  return PCB200_OK;
}

template <int AUXM>
static int launch_rhs_t(const SubstepArgs &a0, int z_lo, int z_hi, size_t part_off, int *nblocks_out)
{
  const pcb200_params &P = g.P;
  SubstepArgs a = a0;
  a.z_lo = z_lo; a.z_hi = z_hi;
  a.partials = g.partials + part_off * NRED;
  const bool M = P.lmagnetic != 0, E = P.lentropy != 0;
  if (z_hi <= z_lo) { *nblocks_out = 0; return PCB200_OK; }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_time_on) { e0 = next_event(); e1 = next_event(); cudaEventRecord(e0, g.st); }
  if (g.use_global) {
    g.kname = "rhs_global";
    dim3 blk(32, 4, 2), grd((P.nx + 31) / 32, (P.ny + 3) / 4, (z_hi - z_lo + 1) / 2);
    *nblocks_out = grd.x * grd.y * grd.z;
    if (M && E) rhs_global<true, true, AUXM><<<grd, blk, 0, g.st>>>(a);
    else if (M) rhs_global<true, false, AUXM><<<grd, blk, 0, g.st>>>(a);
    else if (E) rhs_global<false, true, AUXM><<<grd, blk, 0, g.st>>>(a);
    else rhs_global<false, false, AUXM><<<grd, blk, 0, g.st>>>(a);
  } else if (g.have_tma && !g.use_cpasync) {
    const int ntx_all = (P.nx + TX - 1) / TX, nty_all = (P.ny + TY - 1) / TY;
    const bool sub = a.ntx > 0;                 // tile sub-range (x/y overlap)
    if (!sub) { a.tx0 = 0; a.txs = 1; a.ntx = ntx_all; a.ty0 = 0; a.tys = 1; a.nty = nty_all; }
    a.zchunk = g.zchunk < (z_hi - z_lo) ? g.zchunk : (z_hi - z_lo);
    if (sub) { const int zc = dly_chunk(a.ntx * a.nty, z_hi - z_lo); a.zchunk = zc < a.zchunk ? zc : a.zchunk; }
    dim3 blk(TX, TY), grd(a.ntx, a.nty, (z_hi - z_lo + a.zchunk - 1) / a.zchunk);
    *nblocks_out = grd.x * grd.y * grd.z;
    const CUtensorMap &tm = g.tmap[a.f_in == g.f[0] ? 0 : 1];
    const bool rw = a.alpha != 0.0;
    g.kname = "rhs_tma";
    if (g.use_dly && g.have_dly && P.lbidiagonal_derij && AUXM < 2 && !(E && M && g.me_tma)) {
      const CUtensorMap &tm4 = g.tmap4[a.f_in == g.f[0] ? 0 : 1];
      g.kname = "rhs_dly";
      a.zchunk = sub ? dly_chunk(a.ntx * a.nty, z_hi - z_lo) : g.zchunk_dly;
      if (a.zchunk > z_hi - z_lo) a.zchunk = z_hi - z_lo;
      grd.z = (z_hi - z_lo + a.zchunk - 1) / a.zchunk;
      *nblocks_out = grd.x * grd.y * grd.z;
      // centre-plane-only stencils, z-mixed terms finished three planes late from tensor memory
      constexpr int AX = AUXM < 2 ? AUXM : 0;
      // upwinding is a compile-time property of the kernel (UPW), chosen from the run parameters
      // (the general variant also holds the isentropic eos of runs without the entropy module and gamma /= 1)
      const bool upw = P.lupw_uu || P.lupw_lnrho || (E && P.lupw_ss) || (!E && P.gamma != 1.0);
#define LD(MM, EE, UU) do { if (rw) rhs_dly<MM, EE, 0, true, UU><<<grd, blk, DlyCfg<MM, EE>::SMEM, g.st>>>(a, tm4, g.tmap_wu, g.tmap_wr); \
                            else    rhs_dly<MM, EE, AX, false, UU><<<grd, blk, DlyCfg<MM, EE>::SMEM, g.st>>>(a, tm4, g.tmap_wu, g.tmap_wr); } while (0)
      if (M && E) { if (upw) LD(true, true, true);   else LD(true, true, false); }
      else if (M) { if (upw) LD(true, false, true);  else LD(true, false, false); }
      else if (E) { if (upw) LD(false, true, true);  else LD(false, true, false); }
      else        { if (upw) LD(false, false, true); else LD(false, false, false); }
#undef LD
    } else {
    // the old 2N register is only read on substeps 2,3, which never need the reductions
#define LT(M, E) do { if (P.lbidiagonal_derij) { \
                        if (rw) rhs_tma<M, E, 0, true, true><<<grd, blk, TmaCfg<M, E>::SMEM, g.st>>>(a, tm); \
                        else    rhs_tma<M, E, AUXM, false, true><<<grd, blk, TmaCfg<M, E>::SMEM, g.st>>>(a, tm); \
                      } else { \
                        if (rw) rhs_tma<M, E, 0, true, false><<<grd, blk, TmaCfg<M, E>::SMEM, g.st>>>(a, tm); \
                        else    rhs_tma<M, E, AUXM, false, false><<<grd, blk, TmaCfg<M, E>::SMEM, g.st>>>(a, tm); } } while (0)
    if (M && E) LT(true, true); else if (M) LT(true, false); else if (E) LT(false, true); else LT(false, false);
#undef LT
    }
  } else {
    g.kname = "rhs_march";
    a.zchunk = g.zchunk < (z_hi - z_lo) ? g.zchunk : (z_hi - z_lo);
    dim3 blk(TX, TY), grd((P.nx + TX - 1) / TX, (P.ny + TY - 1) / TY, (z_hi - z_lo + a.zchunk - 1) / a.zchunk);
    *nblocks_out = grd.x * grd.y * grd.z;
    if (M && E) rhs_march<true, true, AUXM><<<grd, blk, MarchCfg<true, true>::SMEM, g.st>>>(a);
    else if (M) rhs_march<true, false, AUXM><<<grd, blk, MarchCfg<true, false>::SMEM, g.st>>>(a);
    else if (E) rhs_march<false, true, AUXM><<<grd, blk, MarchCfg<false, true>::SMEM, g.st>>>(a);
    else rhs_march<false, false, AUXM><<<grd, blk, MarchCfg<false, false>::SMEM, g.st>>>(a);
  }
  if (g_time_on) cudaEventRecord(e1, g.st);
  g.launches++;
  CK(cudaGetLastError());
  if (getenv("PCB200_DEBUG_SYNC")) {      // debugging aid: attribute a device fault to the launch that caused it
    cudaError_t es = cudaStreamSynchronize(g.st);
    if (es != cudaSuccess)
      return fail(PCB200_ECUDA, std::string(g.kname) + " auxm=" + std::to_string(AUXM) + " rw=" + std::to_string(a.alpha != 0.0) + " z=[" +
                  std::to_string(z_lo) + "," + std::to_string(z_hi) + ") zchunk=" + std::to_string(a.zchunk) + ": " + cudaGetErrorString(es));
  }
  return PCB200_OK;
}
static int launch_rhs(int auxm, const SubstepArgs &a, int z_lo, int z_hi, size_t part_off, int *nb)
{
  if (auxm == 0) return launch_rhs_t<0>(a, z_lo, z_hi, part_off, nb);
  if (auxm == 1) return launch_rhs_t<1>(a, z_lo, z_hi, part_off, nb);
  return launch_rhs_t<2>(a, z_lo, z_hi, part_off, nb);
}

// ------------------------------------------------------------------------
// dust particles
static void part_free()
{
  Ctx::Part &p = g.part;
  auto fr = [](auto *&q) { if (q) cudaFree(q); q = nullptr; };
  for (int i = 0; i < 2; i++) { fr(p.fp[i]); fr(p.dfp[i]); fr(p.ipar[i]); fr(p.inear[i]); fr(p.keys[i]); fr(p.vals[i]); fr(p.ckeys[i]); fr(p.cvals[i]); }
  fr(p.imn_array); fr(p.k1); fr(p.k2); fr(p.ccount); fr(p.cs); fr(p.rhop); fr(p.drag); fr(p.aux); fr(p.dt1); fr(p.cub_tmp);
  fr(p.acc); for (int i = 0; i < 4; i++) fr(p.fold[i]);
  fr(p.slot_of_rank); fr(p.counters); fr(p.leave); fr(p.pre); fr(p.holes); fr(p.tails); fr(p.sendbuf); fr(p.recvbuf);
  if (p.counters_host) { cudaFreeHost(p.counters_host); p.counters_host = nullptr; }
  p.fold_elems = 0;
  p.cub_bytes = 0;
  p.on = false;
}

// pencil order of the mn-loop (setup_mm_nn, general.f90:230-335): imn_array[(n-n1)*ny+(m-m1)], 1-based
static void host_imn_array(const pcb200_params &P, std::vector<int> &a)
{
  const int ny = P.ny, nz = P.nz, ncpus = P.nprocx * P.nprocy * P.nprocz;
  a.assign((size_t)ny * nz, 0);
  const int m1 = NG + 1, m2 = NG + ny, n1 = NG + 1, n2 = NG + nz;
  const int m1i = m1 + NG - 1, m2i = m2 - NG + 1, n1i = n1 + NG - 1, n2i = n2 - NG + 1;
  int imn = 1;
  auto visit = [&](int m, int n) { int &v = a[(size_t)(n - n1) * ny + (m - m1)]; if (v == 0) v = imn++; };
  if (ncpus == 1) { for (int n = n1; n <= n2; n++) for (int m = m1; m <= m2; m++) visit(m, n); return; }
  for (int n = n1i + 2; n <= n2i - 2; n++) for (int m = m1i + 2; m <= m2i - 2; m++) visit(m, n);
  for (int n = std::max(n2i - 1, n1 + 1); n <= n2; n++) for (int m = m1i + 2; m <= m2i - 2; m++) visit(m, n);
  for (int n = n1; n <= std::min(n1i + 1, n2); n++) for (int m = m1i + 2; m <= m2i - 2; m++) visit(m, n);
  const int lo = std::min(m1i + 1, m2), hi = std::max(m2i - 1, lo + 1);
  for (int n = n1; n <= n2; n++) { for (int m = m1; m <= lo; m++) visit(m, n); for (int m = hi; m <= m2; m++) visit(m, n); }
}

extern "C" int pcb200_particles_initialize(const pcb200_particle_params *pp)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_particles_initialize before pcb200_initialize");
  if (!pp) return fail(PCB200_EARG, "null particle parameters");
  if (pp->abi_version != PCB200_ABI_VERSION) return fail(PCB200_EARG, "pcb200_particle_params.abi_version mismatch");
  CK(cudaSetDevice(g.dev));
  part_free();
  if (pp->npar_loc <= 0) return PCB200_OK;
  if (pp->mpar_loc < pp->npar_loc) return fail(PCB200_EARG, "mpar_loc < npar_loc");
  if (pp->ldragforce_dust_par && pp->tausp <= 0.0) return fail(PCB200_EARG, "particles: tausp must be > 0");
  const pcb200_params &P = g.P; const DevPar &h = g.hp;
  Ctx::Part &p = g.part;
  p.pp = *pp;
  PartPar &q = p.q; memset(&q, 0, sizeof(q));
  q.npar = pp->npar_loc; q.mpar = pp->mpar_loc;
  q.ldrag_dust = pp->ldragforce_dust_par; q.ldrag_gas = pp->ldragforce_gas_par && pp->ldragforce_dust_par;
  q.lcoriolis = pp->lcoriolis_force_par; q.ldt_adv = pp->ldt_adv_par;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz}, nloc[3] = {P.nx, P.ny, P.nz}, ip[3] = {P.ipx, P.ipy, P.ipz};
  const double hh[3] = {P.dx, P.dy, P.dz}, x0l[3] = {P.x0_loc, P.y0_loc, P.z0_loc};
  for (int a = 0; a < 3; a++) {
    q.wrap[a] = np[a] == 1;            // fold_df/fold_f nproc==1 branches (ghostfold_mpicomm.f90:46,86,120)
    q.lperi[a] = 1;                    // bcpx=bcpy=bcpz='p' (particles_cdata defaults)
    q.shift[a] = nloc[a] * ip[a];
    q.xyz0[a] = P.xyz0[a]; q.Lxyz[a] = P.Lxyz[a]; q.h[a] = hh[a]; q.h_1[a] = 1.0 / hh[a]; q.x0_loc[a] = x0l[a];
  }
  // the index-space formula needs the gas grid's periodicity flag (grid.f90:2330-2334)
  for (int a = 0; a < 3; a++) q.lperi[a] = P.lperi[a] ? 1 : 0;
  q.h_1[0] = P.dx_1; q.h_1[1] = P.dy_1; q.h_1[2] = P.dz_1;
  q.tausp1 = pp->tausp > 0.0 ? 1.0 / pp->tausp : 0.0;
  q.rhop_swarm = pp->rhop_swarm; q.cdtp = pp->cdtp; q.cdtp_drag = pp->cdtp_drag; q.Omega2 = 2.0 * P.Omega;
  q.iux = P.iux - 1; q.ilnrho = P.ilnrho - 1;
  const size_t M = (size_t)pp->mpar_loc;
  for (int i = 0; i < 2; i++) {
    CK(cudaMalloc(&p.fp[i], 6 * M * sizeof(double))); CK(cudaMalloc(&p.dfp[i], 6 * M * sizeof(double)));
    CK(cudaMalloc(&p.ipar[i], M * sizeof(int))); CK(cudaMalloc(&p.inear[i], 3 * M * sizeof(int)));
    CK(cudaMalloc(&p.keys[i], M * sizeof(unsigned))); CK(cudaMalloc(&p.vals[i], M * sizeof(unsigned)));
    CK(cudaMemsetAsync(p.dfp[i], 0, 6 * M * sizeof(double), g.st));
  }
  p.nimn = P.ny * P.nz;
  p.key_bits = 1; while ((1 << p.key_bits) < p.nimn) p.key_bits++;
  std::vector<int> imn; host_imn_array(P, imn);
  CK(cudaMalloc(&p.imn_array, imn.size() * sizeof(int)));
  CK(cudaMemcpyAsync(p.imn_array, imn.data(), imn.size() * sizeof(int), cudaMemcpyHostToDevice, g.st));
  CK(cudaMalloc(&p.k1, p.nimn * sizeof(int))); CK(cudaMalloc(&p.k2, p.nimn * sizeof(int)));
  p.ncell = P.nx * P.ny * P.nz;
  p.cell_bits = 1; while ((1LL << p.cell_bits) < p.ncell) p.cell_bits++;
  for (int i = 0; i < 2; i++) { CK(cudaMalloc(&p.ckeys[i], M * sizeof(unsigned))); CK(cudaMalloc(&p.cvals[i], M * sizeof(unsigned))); }
  CK(cudaMalloc(&p.ccount, (size_t)p.ncell * sizeof(int))); CK(cudaMalloc(&p.cs, ((size_t)p.ncell + 1) * sizeof(int)));
  CK(cudaMalloc(&p.rhop, (size_t)h.mw * sizeof(double))); CK(cudaMalloc(&p.drag, 6 * M * sizeof(double)));   // csort[6][mpar]
  CK(cudaMalloc(&p.aux, (size_t)2 * h.mw * sizeof(double))); CK(cudaMalloc(&p.dt1, sizeof(double)));
  CK(cudaMemsetAsync(p.ccount, 0, (size_t)p.ncell * sizeof(int), g.st));
  CK(cudaMemsetAsync(p.cs, 0, ((size_t)p.ncell + 1) * sizeof(int), g.st));
  CK(cudaMemsetAsync(p.rhop, 0, (size_t)h.mw * sizeof(double), g.st));
  CK(cudaMemsetAsync(p.k1, 0, p.nimn * sizeof(int), g.st)); CK(cudaMemsetAsync(p.k2, 0, p.nimn * sizeof(int), g.st));
  {
    size_t b1 = 0, b2 = 0, b3 = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, b1, p.keys[0], p.keys[1], p.vals[0], p.vals[1], q.mpar, 0, p.key_bits, g.st));
    CK(cub::DeviceRadixSort::SortPairs(nullptr, b2, p.ckeys[0], p.ckeys[1], p.cvals[0], p.cvals[1], q.mpar, 0, p.cell_bits, g.st));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, b3, p.ccount, p.cs, p.ncell, g.st));
    p.cub_bytes = std::max(b1, std::max(b2, b3));
  }
  CK(cudaMalloc(&p.cub_tmp, p.cub_bytes ? p.cub_bytes : 16));
  if (g.nranks > 1) {
    CK(cudaMalloc(&p.acc, (size_t)4 * h.mw * sizeof(double)));
    const size_t pl = std::max((size_t)h.my * h.mz, std::max((size_t)h.mx * h.mz, (size_t)h.mx * h.my)) * 4;
    for (int i = 0; i < 4; i++) CK(cudaMalloc(&p.fold[i], pl * sizeof(double)));
    p.fold_elems = pl;
    // neighbour ranks (26 offsets, periodic, distinct, ascending: the receive order of migrate_particles)
    MigPar &mg = p.mg; memset(&mg, 0, sizeof(mg));
    const int ipq[3] = {P.ipx, P.ipy, P.ipz};
    for (int a = 0; a < 3; a++) { mg.np[a] = np[a]; mg.ip[a] = ipq[a]; mg.nloc[a] = nloc[a]; }
    mg.iproc = g.rank;
    std::vector<int> nb;
    for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) {
      const int qx = (ipq[0] + dx + np[0]) % np[0], qy = (ipq[1] + dy + np[1]) % np[1], qz = (ipq[2] + dz + np[2]) % np[2];
      const int r = qx + np[0] * (qy + np[1] * qz);
      if (r != g.rank && std::find(nb.begin(), nb.end(), r) == nb.end()) nb.push_back(r);
    }
    std::sort(nb.begin(), nb.end());
    mg.nslots = (int)nb.size();
    const char *ce = getenv("PCB200_NPAR_MIG");
    // npar_mig of cparam.local.  Messages have this fixed capacity (the count travels in the last word, so that no
    // host round trip is needed before the exchange): mpar_loc/16 made the exchange of config 5 (1.25e7 particles per
    // rank, 7 neighbours) move 700 MB per substep, 3.2 ms.  mpar_loc/128 is ~3x what a 256^3 block with one particle
    // per cell sends through all six faces at the particle CFL limit; more than that is reported as an error.
    mg.cap = ce && atoi(ce) > 0 ? atoi(ce) : std::max(8192, pp->mpar_loc / 128);
    std::vector<int> sor(g.nranks, -1);
    for (int i = 0; i < mg.nslots; i++) { p.nbr[i] = nb[i]; sor[nb[i]] = i; }
    CK(cudaMalloc(&p.slot_of_rank, g.nranks * sizeof(int)));
    CK(cudaMemcpy(p.slot_of_rank, sor.data(), g.nranks * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&p.counters, 64 * sizeof(int))); CK(cudaMallocHost(&p.counters_host, 64 * sizeof(int)));
    CK(cudaMalloc(&p.leave, M * sizeof(int))); CK(cudaMalloc(&p.pre, (M + 1) * sizeof(int)));
    CK(cudaMalloc(&p.holes, M * sizeof(int))); CK(cudaMalloc(&p.tails, M * sizeof(int)));
    const size_t msg = (size_t)MIG_NVAL * mg.cap + 1;
    CK(cudaMalloc(&p.sendbuf, msg * mg.nslots * sizeof(double))); CK(cudaMalloc(&p.recvbuf, msg * mg.nslots * sizeof(double)));
    size_t b4 = 0, b5 = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, b4, p.leave, p.pre, pp->mpar_loc, g.st));
    CK(cub::DeviceRadixSort::SortPairs(nullptr, b5, p.keys[0], p.keys[1], p.vals[0], p.vals[1], pp->mpar_loc, 0, 5, g.st));
    if (std::max(b4, b5) > p.cub_bytes) { cudaFree(p.cub_tmp); p.cub_bytes = std::max(b4, b5); CK(cudaMalloc(&p.cub_tmp, p.cub_bytes)); }
  }
  CK(cudaMemcpyAsync(p.cs + p.ncell, &q.npar, sizeof(int), cudaMemcpyHostToDevice, g.st));   // after the memset above, same stream
  CK(cudaStreamSynchronize(g.st));
  p.cur = 0; p.on = true;
  return PCB200_OK;
}

extern "C" int pcb200_particles_load(const double *fp, const int32_t *ipar)
{
  Ctx::Part &p = g.part;
  if (!p.on) return fail(PCB200_ESTATE, "pcb200_particles_load: particles are not initialised");
  if (!fp || !ipar) return fail(PCB200_EARG, "null particle arrays");
  const size_t M = (size_t)p.q.mpar;
  CK(cudaMemcpyAsync(p.fp[p.cur], fp, 6 * M * sizeof(double), cudaMemcpyHostToDevice, g.st));
  CK(cudaMemcpyAsync(p.ipar[p.cur], ipar, M * sizeof(int), cudaMemcpyHostToDevice, g.st));
  CK(cudaMemsetAsync(p.dfp[p.cur], 0, 6 * M * sizeof(double), g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

extern "C" int pcb200_particles_store(double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid, int32_t *k1_imn, int32_t *k2_imn)
{
  Ctx::Part &p = g.part;
  if (!p.on) return fail(PCB200_ESTATE, "pcb200_particles_store: particles are not initialised");
  const size_t M = (size_t)p.q.mpar;
  if (fp) CK(cudaMemcpyAsync(fp, p.fp[p.cur], 6 * M * sizeof(double), cudaMemcpyDeviceToHost, g.st));
  if (dfp) CK(cudaMemcpyAsync(dfp, p.dfp[p.cur], 6 * M * sizeof(double), cudaMemcpyDeviceToHost, g.st));
  if (ipar) CK(cudaMemcpyAsync(ipar, p.ipar[p.cur], M * sizeof(int), cudaMemcpyDeviceToHost, g.st));
  if (ineargrid) CK(cudaMemcpyAsync(ineargrid, p.inear[p.cur], 3 * M * sizeof(int), cudaMemcpyDeviceToHost, g.st));
  if (k1_imn) CK(cudaMemcpyAsync(k1_imn, p.k1, p.nimn * sizeof(int), cudaMemcpyDeviceToHost, g.st));
  if (k2_imn) CK(cudaMemcpyAsync(k2_imn, p.k2, p.nimn * sizeof(int), cudaMemcpyDeviceToHost, g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

extern "C" int pcb200_particles_get_aux(double *np, double *rhop)
{
  Ctx::Part &p = g.part;
  if (!p.on) return fail(PCB200_ESTATE, "pcb200_particles_get_aux: particles are not initialised");
  if (!np || !rhop) return fail(PCB200_EARG, "null aux arrays");
  const long long mw = g.hp.mw;
  double *tmp = p.aux;
  part_aux_out<<<(unsigned)((mw + 255) / 256), 256, 0, g.st>>>(p.ccount, p.rhop, p.q.rhop_swarm, tmp, tmp + mw);
  g.launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(np, tmp, mw * sizeof(double), cudaMemcpyDeviceToHost, g.st));
  CK(cudaMemcpyAsync(rhop, tmp + mw, mw * sizeof(double), cudaMemcpyDeviceToHost, g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

extern "C" int pcb200_particles_count(int32_t *npar_loc)
{
  if (!g.part.on) return fail(PCB200_ESTATE, "pcb200_particles_count: particles are not initialised");
  if (npar_loc) *npar_loc = g.part.q.npar;
  return PCB200_OK;
}

// migrate_particles (particles_mpicomm.f90:267-696) after the periodic wrap: leavers out,
// holes refilled from the tail, arrivals appended; one small D2H + sync for the new npar_loc
static int part_migrate()
{
  Ctx::Part &p = g.part; PartPar &q = p.q; const MigPar &mg = p.mg;
  const unsigned nb = (unsigned)std::max(1, (q.npar + 255) / 256);
  CK(cudaMemsetAsync(p.counters, 0, 64 * sizeof(int), g.st));
  part_mig_classify<<<nb, 256, 0, g.st>>>(q, mg, p.fp[p.cur], p.slot_of_rank, p.keys[0], p.vals[0], p.leave, p.counters);
  part_mig_offsets<<<1, 32, 0, g.st>>>(mg, q.npar, p.counters);
  size_t bytes = p.cub_bytes;
  CK(cub::DeviceScan::ExclusiveSum(p.cub_tmp, bytes, p.leave, p.pre, q.npar, g.st));
  bytes = p.cub_bytes;
  CK(cub::DeviceRadixSort::SortPairs(p.cub_tmp, bytes, p.keys[0], p.keys[1], p.vals[0], p.vals[1], q.npar, 0, 5, g.st));
  const unsigned nbp = (unsigned)std::max(1, (std::max(q.npar, mg.nslots) + 255) / 256);
  part_mig_pack<<<nbp, 256, 0, g.st>>>(mg, q.npar, q.mpar, p.keys[1], p.vals[1], p.fp[p.cur], p.dfp[p.cur], p.ipar[p.cur], p.counters, p.sendbuf);
  part_mig_holes<<<nb, 256, 0, g.st>>>(q.npar, p.leave, p.pre, p.counters, p.holes, p.tails);
  part_mig_fill<<<nb, 256, 0, g.st>>>(q.npar, q.mpar, p.pre, p.counters, p.holes, p.tails, p.fp[p.cur], p.dfp[p.cur], p.ipar[p.cur]);
  g.launches += 5;
  CK(cudaGetLastError());
  ghost_mark(PM_MIG_K, g.st);
  const size_t msg = (size_t)MIG_NVAL * mg.cap + 1;
  NK(g.nccl.GroupStart());
  for (int s = 0; s < mg.nslots; s++) {
    NK(g.nccl.Send(p.sendbuf + (size_t)s * msg, msg, 8, p.nbr[s], g.comm, g.st));
    NK(g.nccl.Recv(p.recvbuf + (size_t)s * msg, msg, 8, p.nbr[s], g.comm, g.st));
  }
  NK(g.nccl.GroupEnd());
  dim3 grd((mg.cap + 255) / 256, mg.nslots);
  part_mig_append<<<grd, 256, 0, g.st>>>(mg, q.mpar, p.recvbuf, p.counters, p.fp[p.cur], p.dfp[p.cur], p.ipar[p.cur]);
  g.launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(p.counters_host, p.counters, 64 * sizeof(int), cudaMemcpyDeviceToHost, g.st));
  CK(cudaStreamSynchronize(g.st));
  const int err = p.counters_host[29];
  if (err & 1) return fail(PCB200_EARG, "migrate_particles: a particle wants to migrate to a processor that is not a neighbour");
  if (err & 2) return fail(PCB200_EARG, "migrate_particles: too many particles migrating (raise PCB200_NPAR_MIG)");
  if (err & 4) return fail(PCB200_EARG, "migrate_particles: too many particles want to be at this rank (mpar_loc too small)");
  q.npar = p.counters_host[30];
  ghost_mark(PM_MIG_X, g.st);
  return PCB200_OK;
}

// particles_timestep_first + particles_boundconds (timestep.f90:127, equ.f90:169)
static int part_before_rhs(int isub)
{
  Ctx::Part &p = g.part; PartPar &q = p.q; const DevPar &h = g.hp;
  unsigned nb = (unsigned)std::max(1, (q.npar + 255) / 256);
  part_first<<<nb, 256, 0, g.st>>>(p.dfp[p.cur], q.npar, q.mpar, isub == 1 ? 0.0 : g.P.alpha_ts[isub - 1]);
  g.launches++;
  if (g.nranks > 1) {
    part_wrap<<<nb, 256, 0, g.st>>>(q, p.fp[p.cur]);
    g.launches++;
    ghost_mark(PM_FIRST, g.st);
    int rcm = part_migrate();
    if (rcm) return rcm;
    nb = (unsigned)std::max(1, (q.npar + 255) / 256);
  }
  part_bc_map<<<nb, 256, 0, g.st>>>(q, p.fp[p.cur], p.inear[p.cur], p.imn_array, p.keys[0], p.vals[0]);
  g.launches++;
  CK(cudaGetLastError());
  size_t bytes = p.cub_bytes;
  CK(cub::DeviceRadixSort::SortPairs(p.cub_tmp, bytes, p.keys[0], p.keys[1], p.vals[0], p.vals[1], q.npar, 0, p.key_bits, g.st));
  const int nx = 1 - p.cur;
  part_gather<<<nb, 256, 0, g.st>>>(q.npar, q.mpar, p.vals[1], p.fp[p.cur], p.fp[nx], p.dfp[p.cur], p.dfp[nx],
                                    p.ipar[p.cur], p.ipar[nx], p.inear[p.cur], p.inear[nx]);
  p.cur = nx;
  CK(cudaMemsetAsync(p.k1, 0, p.nimn * sizeof(int), g.st)); CK(cudaMemsetAsync(p.k2, 0, p.nimn * sizeof(int), g.st));
  part_pencil_bounds<<<nb, 256, 0, g.st>>>(q.npar, p.keys[1], p.k1, p.k2);
  ghost_mark(PM_MAP, g.st);
  // cell bins for the gather-deposit: count per cell (= np), particles sorted by cell, bin offsets
  CK(cudaMemsetAsync(p.ccount, 0, ((size_t)p.ncell) * sizeof(int), g.st));
  part_cellkey<<<nb, 256, 0, g.st>>>(q.npar, q.mpar, p.inear[p.cur], p.ckeys[0], p.cvals[0], p.ccount);
  g.launches += 2;
  CK(cudaGetLastError());
  bytes = p.cub_bytes;
  CK(cub::DeviceRadixSort::SortPairs(p.cub_tmp, bytes, p.ckeys[0], p.ckeys[1], p.cvals[0], p.cvals[1], q.npar, 0, p.cell_bits, g.st));
  part_invperm<<<nb, 256, 0, g.st>>>(q.npar, p.cvals[1], p.cvals[0]);      // cvals[0] := rank of each particle in its bin list
  g.launches++;
  bytes = p.cub_bytes;
  // cs[ncell] = npar (npar changes when particles migrate)
  if (p.counters) CK(cudaMemcpyAsync(p.cs + p.ncell, p.counters + 30, sizeof(int), cudaMemcpyDeviceToDevice, g.st));
  CK(cub::DeviceScan::ExclusiveSum(p.cub_tmp, bytes, p.ccount, p.cs, p.ncell, g.st));
  ghost_mark(PM_CELLS, g.st);
  (void)h;
  return PCB200_OK;
}

// the particle share of the mn-loop (equ.f90:1046) + dxxp_dt/dvvp_dt (:379-382): needs valid ghost zones
static int part_pencils(const double *fin, int ldt)
{
  Ctx::Part &p = g.part; const PartPar &q = p.q; const pcb200_params &P = g.P;
  if (ldt) CK(cudaMemsetAsync(p.dt1, 0, sizeof(double), g.st));
  part_drag<<<(unsigned)std::max(1, (q.npar + 127) / 128), 128, 0, g.st>>>(q, fin, p.fp[p.cur], p.dfp[p.cur], p.inear[p.cur], p.cvals[0], p.drag, p.dt1, ldt);
  g.launches++;
  if (ldt && q.ldrag_dust) {
    dim3 blk(128), grd((P.nx + 127) / 128, P.ny, P.nz);
    part_dt_cells<<<grd, blk, 0, g.st>>>(q, fin, p.ccount, p.dt1);
    g.launches++;
  }
  CK(cudaGetLastError());
  return PCB200_OK;
}

// fold_df/fold_f between ranks (ghostfold_mpicomm.f90:28-153): the first ghost layer of the
// deposit sums goes to the neighbour rank and is added to its first interior plane; z, then
// y, then x; the folded layer is zeroed.  Neighbours are periodic whatever the gas boundary
// condition is (zlneigh/zuneigh of mpicomm.f90).  acc: [4][mw].
static int part_fold_exchange(double *acc)
{
  Ctx::Part &p = g.part; const pcb200_params &P = g.P; const DevPar &h = g.hp;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz}, nloc[3] = {P.nx, P.ny, P.nz};
  const int ip[3] = {P.ipx, P.ipy, P.ipz};
  for (int axis = 2; axis >= 0; axis--) {
    if (np[axis] == 1) continue;
    auto rank_of = [&](int i) { int q[3] = {ip[0], ip[1], ip[2]}; q[axis] = (i + np[axis]) % np[axis]; return q[0] + np[0] * (q[1] + np[1] * q[2]); };
    // axis communicator: ranks are positions along the axis
    ncclComm_t cm = g.comm_ax[axis] ? g.comm_ax[axis] : g.comm;
    const int nlo = g.comm_ax[axis] ? (ip[axis] + np[axis] - 1) % np[axis] : rank_of(ip[axis] - 1);
    const int nhi = g.comm_ax[axis] ? (ip[axis] + 1) % np[axis] : rank_of(ip[axis] + 1);
    const long long per = axis == 0 ? (long long)h.my * h.mz : (axis == 1 ? (long long)h.mx * h.mz : (long long)h.mx * h.my);
    const size_t cnt = (size_t)per * 4;
    dim3 blk(256), grd((unsigned)((per + 255) / 256), 4);
    auto copy = [&](double *buf, int plane, int mode) {
      FaceArgs a; a.f = acc; a.buf = buf; a.axis = axis; a.start = plane; a.unpack = mode; a.nplanes = 1;
      face_copy<<<grd, blk, 0, g.st>>>(a);
      g.launches++;
    };
    const int n = nloc[axis];
    copy(p.fold[0], NG - 1, 0);           // my lower ghost layer -> lower neighbour's top interior plane
    copy(p.fold[1], NG + n, 0);           // my upper ghost layer -> upper neighbour's bottom interior plane
    NK(g.nccl.GroupStart());
    NK(g.nccl.Send(p.fold[0], cnt, 8, nlo, cm, g.st));
    NK(g.nccl.Recv(p.fold[3], cnt, 8, nhi, cm, g.st));
    NK(g.nccl.Send(p.fold[1], cnt, 8, nhi, cm, g.st));
    NK(g.nccl.Recv(p.fold[2], cnt, 8, nlo, cm, g.st));
    NK(g.nccl.GroupEnd());
    copy(p.fold[3], NG + n - 1, 2);       // += upper neighbour's lower ghost layer
    copy(p.fold[2], NG, 2);               // += lower neighbour's upper ghost layer
    plane_zero<<<grd, blk, 0, g.st>>>(acc, axis, NG - 1);
    plane_zero<<<grd, blk, 0, g.st>>>(acc, axis, NG + n);
    g.launches += 2;
    CK(cudaGetLastError());
  }
  return PCB200_OK;
}

// PCB200_TIME_GHOST=1: CUDA events around the ghost-zone section (boundary conditions, packed x/y faces,
// z planes) of every substep on the compute stream; the sum is printed by pcb200_finalize
static int do_substep(int isub, int ldiagnos)
{
  const pcb200_params &P = g.P;
  double *fin = g.f[g.cur], *fout = g.f[1 - g.cur];
  int rc;
  // Ghost zones, axis by axis in the reference order x -> y -> z (equ.f90:211-229,
  // boundcond.f90:1283-1290), each over the full extent of the other axes.  Per axis: wrap
  // locally / physical boundary condition / exchange with the neighbour ranks.  Consecutive
  // locally periodic axes share one launch.  Only the z exchange overlaps with compute.
  ghost_mark(GM_START, g.st);
  if (g.part.on) if ((rc = part_before_rhs(isub))) return rc;
  const int np[3] = {P.nprocx, P.nprocy, P.nprocz};
  if (g.part.on) ghost_mark(PM_CELLS, g.st);
  bool pend[3] = {false, false, false};
  auto flush = [&]() -> int {
    if (!pend[0] && !pend[1] && !pend[2]) return PCB200_OK;
    const int r = launch_bc_periodic(fin, pend[0], pend[1], pend[2]);
    pend[0] = pend[1] = pend[2] = false;
    return r;
  };
  const bool multi = np[2] > 1;          // z decomposed
  const bool first = isub == 1;
  const bool need_dt = first && P.ldt;
  // Isothermal runs: dt from a light CFL pre-pass over the start-of-step state, so that substep 1
  // applies the 2N update itself and the separate axpy pass over f and w disappears.
  // (isothermal = no entropy module and gamma = 1; with gamma /= 1 the sound speed varies from cell to cell)
  const bool prepass = need_dt && !ldiagnos && !P.lentropy && P.gamma == 1.0 && g.use_prepass;
  // x/y-decomposed runs: the whole ghost-zone sequence (boundary conditions, packed x/y faces, z
  // planes) runs on the comm stream while the compute stream works on the tiles that need no ghost
  // cell; the boundary shell (first/last tile column and row, first/last NG planes) follows
  // (initiate_isendrcv_bdry / finalize_isendrcv_bdry, mpicomm.f90:1325-2047).  Not in the substep
  // that starts with the CFL pre-pass (it needs the ghost zones at once).
  const int ntx = (P.nx + TX - 1) / TX, nty = (P.ny + TY - 1) / TY;
  // Off by default: measured slower than the sequential order (2 GPUs split in x: 29.3 vs 28.4 ms/step; 8 GPUs
  // 2x2x2: 35.5 vs 34.3 ms/step) -- the compute kernel holds every SM with one 1-ms CTA each, so the small
  // kernels of the comm stream do not start on time.  PCB200_OVERLAP_XY=1 switches it on.
  const bool ovl_want = g.overlap_force > 0;
  const bool ovl = (np[0] > 1 || np[1] > 1) && !g.part.on && !prepass && g.have_tma && !g.use_global &&
                   !g.use_cpasync && ntx >= 3 && nty >= 3 && P.nz >= 3 * NG + 2 && ovl_want;
  if (ovl) {
    CK(cudaEventRecord(g.ev_bc, g.st));
    CK(cudaStreamWaitEvent(g.st_comm, g.ev_bc, 0));
    g.bcs = g.st_comm;
  }
  for (int axis = 0; axis < 3; axis++) {
    if (np[axis] == 1) {
      if (wraps_locally(axis)) { pend[axis] = true; continue; }
      if ((rc = flush())) return rc;
      if ((rc = launch_bc(fin, axis, 1, 1))) return rc;
      continue;
    }
    if ((rc = flush())) return rc;
    int nlo, nhi; bool lfirst, llast;
    neighbours(axis, &nlo, &nhi, &lfirst, &llast);
    const int lo = lfirst && nlo < 0, hi = llast && nhi < 0;     // physical boundaries (lfirst_proc_*/llast_proc_*)
    if (axis < 2) {
      if (lo || hi) if ((rc = launch_bc(fin, axis, lo, hi))) return rc;
      if ((rc = exchange_packed(fin, axis))) return rc;
    } else {
      if (!ovl) { CK(cudaEventRecord(g.ev_bc, g.st)); CK(cudaStreamWaitEvent(g.st_comm, g.ev_bc, 0)); }
      ghost_mark(GM_BC, g.st);
      if ((rc = exchange_z(fin))) return rc;
      ghost_mark(GM_Z, g.st_comm);
      if (!ovl) CK(cudaEventRecord(g.ev_halo, g.st_comm));
      if (lo || hi) if ((rc = launch_bc(fin, 2, lo, hi))) return rc;
    }
  }
  if ((rc = flush())) return rc;
  if (ovl) { CK(cudaEventRecord(g.ev_halo, g.st_comm)); g.bcs = g.st; }
  if (g_time_ghost) {
    ghost_mark(GM_BC, g.st);
    if (multi && !ovl) { CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0)); ghost_mark(GM_END, g.st); }
  }
  const int auxm = first ? (ldiagnos ? 2 : (need_dt ? 1 : 0)) : 0;
  SubstepArgs a; memset(&a, 0, sizeof(a));
  a.f_in = fin; a.w = g.w; a.f_out = fout;
  a.alpha = first ? 0.0 : P.alpha_ts[isub - 1];
  a.beta = P.beta_ts[isub - 1];
  a.dt_dev = g.dt_dev;
  a.write_w = isub < 3;
  a.write_f = need_dt ? 0 : 1;
  if (need_dt) a.write_w = 1;
  if (g.part.on) {
    if (multi) { CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0)); }
    if ((rc = part_pencils(fin, need_dt))) return rc;
    ghost_mark(PM_DRAG, g.st);
  }
  int nb0 = 0, nb1 = 0, nb2 = 0;
  if (ovl) {
    int nb = 0; size_t off = 0;
    // Boundary conditions 'a' and 'cT' also set the boundary plane itself (an interior cell), so the
    // overlapped part keeps one more plane away from the z faces: it reads z-3 >= 1.
    const int ZB = NG + 1;
    SubstepArgs t = a;                   // tiles that touch no ghost cell
    t.tx0 = 1; t.txs = 1; t.ntx = ntx - 2; t.ty0 = 1; t.tys = 1; t.nty = nty - 2;
    if ((rc = launch_rhs(auxm, t, ZB, P.nz - ZB, off, &nb))) return rc;
    off += nb;
    CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0));
    if ((rc = launch_rhs(auxm, a, 0, ZB, off, &nb))) return rc;                  // first and last ZB planes, all tiles
    off += nb;
    if ((rc = launch_rhs(auxm, a, P.nz - ZB, P.nz, off, &nb))) return rc;
    off += nb;
    t = a; t.tx0 = 0; t.txs = 1; t.ntx = ntx; t.ty0 = 0; t.tys = nty - 1; t.nty = 2;   // first and last tile row
    if ((rc = launch_rhs(auxm, t, ZB, P.nz - ZB, off, &nb))) return rc;
    off += nb;
    t = a; t.tx0 = 0; t.txs = ntx - 1; t.ntx = 2; t.ty0 = 1; t.tys = 1; t.nty = nty - 2; // first and last tile column
    if ((rc = launch_rhs(auxm, t, ZB, P.nz - ZB, off, &nb))) return rc;
    off += nb;
    nb0 = (int)off;
  } else if (prepass) {
    if (multi) CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0));
    const char *zce = getenv("PCB200_PREPASS_ZC");
    const int zc = zce && atoi(zce) >= 16 ? atoi(zce) : 64;      // g.partials is sized for chunks of >= 16 planes
    dim3 blk(TX, TY), grd((P.nx + TX - 1) / TX, (P.ny + TY - 1) / TY, (P.nz + zc - 1) / zc);
    if (P.lmagnetic) cfl_prepass<true><<<grd, blk, 0, g.st>>>(fin, g.partials, 0, P.nz, zc);
    else cfl_prepass<false><<<grd, blk, 0, g.st>>>(fin, g.partials, 0, P.nz, zc);
    reduce_partials<<<1, NRED * 32, 0, g.st>>>(g.partials, (int)(grd.x * grd.y * grd.z), g.red_dev);
    const double floor1 = P.dtmax != 0.0 ? 1.0 / P.dtmax : 0.0;
    finish_dt<<<1, 32, 0, g.st>>>(g.red_dev, floor1, g.dt_dev, g.part.on ? g.part.dt1 : nullptr);
    g.launches += 3;
    if (g.nranks > 1) {
      NK(g.nccl.AllReduce(g.dt_dev + 1, g.dt_dev + 1, 1, 8, 2 /*ncclMax*/, g.comm, g.st));     // set_dt, sub.f90:3909
      set_dt_from_dt1<<<1, 32, 0, g.st>>>(g.dt_dev);
      g.launches++;
    }
    CK(cudaGetLastError());
    a.write_f = 1;
    if ((rc = launch_rhs(0, a, 0, P.nz, 0, &nb0))) return rc;
  } else if (!multi) {
    if ((rc = launch_rhs(auxm, a, 0, P.nz, 0, &nb0))) return rc;
  } else if (!g.zsplit) {
    // default: wait for the z halo (~0.2 ms), then one launch (no interior/boundary split)
    CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0));
    if ((rc = launch_rhs(auxm, a, 0, P.nz, 0, &nb0))) return rc;
  } else {
    // interior planes overlap with the halo exchange (general.f90:267-279 / gpu_astaroth.cc:162-166)
    if ((rc = launch_rhs(auxm, a, NG, P.nz - NG, 0, &nb0))) return rc;
    CK(cudaStreamWaitEvent(g.st, g.ev_halo, 0));
    if ((rc = launch_rhs(auxm, a, 0, NG, nb0, &nb1))) return rc;
    if ((rc = launch_rhs(auxm, a, P.nz - NG, P.nz, (size_t)nb0 + nb1, &nb2))) return rc;
  }
  if (g.part.on) {
    // TSC gather: rhop, and the drag back-reaction joins the 2N register and (when this launch
    // already produced the new state) f_out; otherwise axpy_update carries it (equ.f90:1046)
    Ctx::Part &pt = g.part;
    dim3 blk(128), grd((P.nx + 127) / 128, P.ny, P.nz);
    ghost_mark(PM_GAS, g.st);
    if (g.nranks == 1) {
      part_deposit<false><<<grd, blk, 0, g.st>>>(pt.q, fin, pt.drag, pt.cs, pt.rhop, g.w, a.write_f ? fout : nullptr, g.dt_dev, a.beta, nullptr);
      g.launches++;
      ghost_mark(PM_DEPOSIT, g.st);
    } else {
      // split directions deposit into the first ghost layer, which fold_df ships to the neighbour
      CK(cudaMemsetAsync(pt.acc, 0, (size_t)4 * g.hp.mw * sizeof(double), g.st));
      dim3 grg((P.nx + 2 + 127) / 128, P.ny + 2, P.nz + 2);
      part_deposit<true><<<grg, blk, 0, g.st>>>(pt.q, fin, pt.drag, pt.cs, nullptr, nullptr, nullptr, g.dt_dev, 0.0, pt.acc);
      g.launches++;
      ghost_mark(PM_DEPOSIT, g.st);
      if ((rc = part_fold_exchange(pt.acc))) return rc;
      ghost_mark(PM_FOLD, g.st);
      part_backreact_acc<<<grd, blk, 0, g.st>>>(pt.q, fin, pt.acc, pt.rhop, g.w, a.write_f ? fout : nullptr, g.dt_dev, a.beta);
      g.launches++;
      ghost_mark(PM_BACK, g.st);
    }
    CK(cudaGetLastError());
  }
  if (auxm > 0 && !prepass) {
    reduce_partials<<<1, NRED * 32, 0, g.st>>>(g.partials, nb0 + nb1 + nb2, g.red_dev);
    g.launches++;
    CK(cudaGetLastError());
  }
  if (need_dt && !prepass) {
    const double floor1 = P.dtmax != 0.0 ? 1.0 / P.dtmax : 0.0;
    finish_dt<<<1, 32, 0, g.st>>>(g.red_dev, floor1, g.dt_dev, g.part.on ? g.part.dt1 : nullptr);
    g.launches++;
    if (g.nranks > 1) {
      NK(g.nccl.AllReduce(g.dt_dev + 1, g.dt_dev + 1, 1, 8, 2 /*ncclMax*/, g.comm, g.st));     // set_dt, sub.f90:3909
      set_dt_from_dt1<<<1, 32, 0, g.st>>>(g.dt_dev);
      g.launches++;
    }
    dim3 blk(128), grd((P.nx + 127) / 128, P.ny, P.nz * P.mvar);
    axpy_update<<<grd, blk, 0, g.st>>>(fin, g.w, fout, g.dt_dev, P.beta_ts[0]);
    g.launches++;
    CK(cudaGetLastError());
  }
  if (g.part.on) {
    part_second<<<(unsigned)std::max(1, (g.part.q.npar + 255) / 256), 256, 0, g.st>>>(g.part.fp[g.part.cur], g.part.dfp[g.part.cur],
                                                                         g.part.q.npar, g.part.q.mpar, g.dt_dev, P.beta_ts[isub - 1]);
    g.launches++;
    CK(cudaGetLastError());
  }
  g.cur ^= 1;
  if (isub == 3 && g.have_forcing) {
    ForcingArgs fa;
    for (int j = 0; j < 3; j++) { fa.coef1[j] = g.frc.coef1[j]; fa.coef2[j] = g.frc.coef2[j]; fa.fda[j] = g.frc.fda[j]; fa.kk[j] = g.frc.kk[j]; }
    fa.phase = g.frc.phase; fa.fact = g.frc.fact;
    fa.x0 = P.x0_loc; fa.y0 = P.y0_loc; fa.z0 = P.z0_loc; fa.dx = P.dx; fa.dy = P.dy; fa.dz = P.dz;
    dim3 blk(128), grd((P.nx + 127) / 128, P.ny, P.nz);
    forcing_kick<<<grd, blk, 0, g.st>>>(g.f[g.cur], fa);
    g.launches++;
    CK(cudaGetLastError());
  }
  if (first && (need_dt || ldiagnos)) {
    // dt (and the diagnostics) go back to the host: one small D2H + sync per time step
    CK(cudaMemcpyAsync(g.pinned, g.dt_dev, sizeof(double) * 2, cudaMemcpyDeviceToHost, g.st));
    if (auxm > 0) CK(cudaMemcpyAsync(g.pinned + 2, g.red_dev, sizeof(double) * NRED, cudaMemcpyDeviceToHost, g.st));
    CK(cudaStreamSynchronize(g.st));
    if (need_dt) g.dt = g.pinned[0];
    if (ldiagnos) {
      const double *r = g.pinned + 2;
      double *d = g.diag;
      d[PCB200_DIAG_U2SUM] = r[RS_U2]; d[PCB200_DIAG_U2MAX] = r[RD_U2MAX]; d[PCB200_DIAG_RHOSUM] = r[RS_RHO];
      d[PCB200_DIAG_SSSUM] = r[RS_SS]; d[PCB200_DIAG_B2SUM] = r[RS_B2]; d[PCB200_DIAG_B2MAX] = r[RD_B2MAX];
      d[PCB200_DIAG_J2SUM] = r[RS_J2]; d[PCB200_DIAG_J2MAX] = r[RD_J2MAX]; d[PCB200_DIAG_DTC] = r[RD_DTC];
      d[PCB200_DIAG_DTU] = r[RD_DTU]; d[PCB200_DIAG_DTNU] = r[RD_DTNU]; d[PCB200_DIAG_DTCHI] = r[RD_DTCHI];
      d[PCB200_DIAG_DT1MAX] = need_dt ? fmax(r[RD_DT1], P.dtmax != 0.0 ? 1.0 / P.dtmax : 0.0) : r[RD_DT1];
      d[PCB200_DIAG_ABSUM] = r[RS_AB]; d[PCB200_DIAG_JBSUM] = r[RS_JB]; d[PCB200_DIAG_OUSUM] = r[RS_OU];
    }
  }
  return PCB200_OK;
}

extern "C" int pcb200_substep(int isubstep, int full, int early_finalize, int ldiagnos, double *dt)
{
  (void)early_finalize;
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_substep before pcb200_initialize");
  if (isubstep < 1 || isubstep > 3) return fail(PCB200_EARG, "isubstep must be 1..3");
  CK(cudaSetDevice(g.dev));
  int rc;
  if (full) { if ((rc = pcb200_load_farray(nullptr))) return rc; }
  if (isubstep == 1 && !g.P.ldt) {
    // fixed time step: the caller's dt (or dt_fixed) is authoritative
    g.dt = (dt && *dt > 0.0) ? *dt : g.P.dt_fixed;
    if (!(g.dt > 0.0)) return fail(PCB200_EARG, "pcb200_substep: ldt=0 needs a positive time step (*dt or dt_fixed)");
    g.pinned[0] = g.dt; g.pinned[1] = 1.0 / g.dt;
    CK(cudaMemcpyAsync(g.dt_dev, g.pinned, sizeof(double) * 2, cudaMemcpyHostToDevice, g.st));
    CK(cudaStreamSynchronize(g.st));
  }
  if ((rc = do_substep(isubstep, ldiagnos))) return rc;
  if (dt) *dt = g.dt;
  return PCB200_OK;
}

extern "C" int pcb200_timestep(int full, int ldiagnos, double *dt)
{
  int rc;
  for (int s = 1; s <= 3; s++) if ((rc = pcb200_substep(s, s == 1 ? full : 0, 0, s == 1 ? ldiagnos : 0, dt))) return rc;
  return PCB200_OK;
}

extern "C" int pcb200_copy_farray(double *f_host)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_copy_farray before pcb200_initialize");
  double *dst = f_host ? f_host : g.f_host;
  if (!dst) return fail(PCB200_ESTATE, "no host array registered");
  CK(cudaSetDevice(g.dev));
  if (dst == g.f_host) pin_registered();
  CK(cudaMemcpyAsync(dst, g.f[g.cur], g.nelem * sizeof(double), cudaMemcpyDeviceToHost, g.st));
  CK(cudaStreamSynchronize(g.st));
  return PCB200_OK;
}

// wsnap (snapshot.f90:286-302) + output_snap (io_dist.f90:110-167) from the device, overlapped with the next steps
extern "C" int pcb200_snapshot_begin(const char *path, double t, const double *x, const double *y, const double *z,
                                     double dx, double dy, double dz)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_snapshot_begin before pcb200_initialize");
  if (!path || !x || !y || !z) return fail(PCB200_EARG, "pcb200_snapshot_begin: null argument");
  CK(cudaSetDevice(g.dev));
  SnapWriter &w = g_snap;
  if (w.wait()) return fail(w.rc, w.err);                  // one snapshot in flight at a time
  w.device = g.dev;
  if (w.nelem != g.nelem) { if (w.dev) cudaFree(w.dev); w.dev = nullptr; CK(cudaMalloc(&w.dev, g.nelem * sizeof(double))); w.nelem = g.nelem; }
  if (!w.st) {
    CK(cudaStreamCreateWithFlags(&w.st, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&w.ev_copy, cudaEventDisableTiming));
    for (int b = 0; b < 2; b++) CK(cudaEventCreateWithFlags(&w.ev_chunk[b], cudaEventDisableTiming));
  }
  if (!w.pin[0]) {
    const char *ce = getenv("PCB200_SNAP_CHUNK_MB");
    w.chunk = (size_t)(ce && atoi(ce) > 0 ? atoi(ce) : 64) << 20;
    for (int b = 0; b < 2; b++) CK(cudaMallocHost(&w.pin[b], w.chunk));
  }
  // the only work on the compute stream: f -> staging array (ordered after the last substep)
  CK(cudaMemcpyAsync(w.dev, g.f[g.cur], g.nelem * sizeof(double), cudaMemcpyDeviceToDevice, g.st));
  CK(cudaEventRecord(w.ev_copy, g.st));
  const DevPar &h = g.hp;
  std::vector<double> tail;
  tail.reserve(4 + h.mx + h.my + h.mz);
  tail.push_back(t);
  tail.insert(tail.end(), x, x + h.mx); tail.insert(tail.end(), y, y + h.my); tail.insert(tail.end(), z, z + h.mz);
  tail.push_back(dx); tail.push_back(dy); tail.push_back(dz);
  const char *re = getenv("PCB200_SNAP_RECMAX");              // tests: exercise the subrecord split with small files
  const long long rec_max = re && atoll(re) > 0 ? atoll(re) : 2147483639LL;
  w.busy = true;
  w.th = std::thread(&SnapWriter::run, &w, std::string(path), std::move(tail), rec_max);
  return PCB200_OK;
}

extern "C" int pcb200_snapshot_wait(void)
{
  if (g_snap.wait()) return fail(g_snap.rc, g_snap.err);
  return PCB200_OK;
}

extern "C" int pcb200_get_diagnostics(double *out, int n)
{
  if (!out || n < 1) return fail(PCB200_EARG, "bad diagnostics buffer");
  for (int i = 0; i < n && i < PCB200_NDIAG; i++) out[i] = g.diag[i];
  return PCB200_OK;
}

extern "C" int pcb200_nccl_unique_id(char id128[128])
{
  if (!g.nccl.h) {
    const char *names[] = {"libnccl.so.2", "libnccl.so", nullptr};
    for (int i = 0; names[i] && !g.nccl.h; i++) g.nccl.h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!g.nccl.h) return fail(PCB200_ENCCL, std::string("cannot dlopen libnccl: ") + dlerror());
#define SYM(n) *(void **)(&g.nccl.n) = dlsym(g.nccl.h, "nccl" #n); if (!g.nccl.n) return fail(PCB200_ENCCL, "missing symbol nccl" #n)
    SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(Send); SYM(Recv); SYM(AllReduce); SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
#undef SYM
    *(void **)(&g.nccl.CommSplit) = dlsym(g.nccl.h, "ncclCommSplit");     // absent before 2.18: the exchanges then use the parent communicator
  }
  if (id128) { ncclUniqueId id; NK(g.nccl.GetUniqueId(&id)); memcpy(id128, id.internal, 128); }
  return PCB200_OK;
}

extern "C" int pcb200_comm_init(const char id128[128], int rank, int nranks)
{
  if (!g.inited) return fail(PCB200_ESTATE, "pcb200_comm_init before pcb200_init");
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail(PCB200_EARG, "bad rank/nranks");
  g.rank = rank; g.nranks = nranks;
  if (nranks == 1) return PCB200_OK;
  int rc = pcb200_nccl_unique_id(nullptr);
  if (rc) return rc;
  CK(cudaSetDevice(g.dev));
  ncclUniqueId id; memcpy(id.internal, id128, 128);
  NK(g.nccl.CommInitRank(&g.comm, nranks, id, rank));
  return PCB200_OK;
}

extern "C" int pcb200_finalize(void)
{
  if (!g.inited) return PCB200_OK;
  cudaSetDevice(g.dev);
  cudaStreamSynchronize(g.st);
  if (g_time_ghost && !g_ghost_ev.empty()) {
    // the first two time steps are left out (ranks start skewed: the first exchange waits for the slowest set-up)
    double tot[GM_N] = {0}; size_t nsub = 0, seen = 0;
    for (size_t i = 0; i < g_ghost_ev.size(); i++) {
      if (g_ghost_ev[i].first == GM_START) { seen++; if (seen > 6) nsub++; continue; }
      if (seen <= 6) continue;
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, g_ghost_ev[i - 1].second, g_ghost_ev[i].second) == cudaSuccess) tot[g_ghost_ev[i].first] += ms;
    }
    fprintf(stderr, "[pcb200 rank %d] ghost-zone section, ms per substep over %zu substeps:", g.rank, nsub);
    double all = 0.0;
    for (int t = 1; t < GM_N; t++) { fprintf(stderr, " %s %.3f", g_gm_name[t], nsub ? tot[t] / nsub : 0.0); all += tot[t]; }
    fprintf(stderr, " | total %.3f\n", nsub ? all / nsub : 0.0);
    for (auto &e : g_ghost_ev) cudaEventDestroy(e.second);
    g_ghost_ev.clear();
  }
  free_fields();
  if (g.f_host_pinned) { cudaHostUnregister(g.f_host); g.f_host_pinned = false; }
  g.f_host = nullptr;
  for (int q = 0; q < PCB200_NPROF; q++) { if (g.zprof[q]) cudaFree(g.zprof[q]); g.zprof[q] = nullptr; }
  for (int i = 0; i < 4; i++) { if (g.face[i]) cudaFree(g.face[i]); g.face[i] = nullptr; }
  g.face_elems = 0;
  part_free();
  for (int a = 0; a < 3; a++) { if (g.comm_ax[a] && g.nccl.CommDestroy) g.nccl.CommDestroy(g.comm_ax[a]); g.comm_ax[a] = nullptr; g.split_np[a] = 0; }
  if (g.comm && g.nccl.CommDestroy) { g.nccl.CommDestroy(g.comm); g.comm = nullptr; }
  g.have_par = false; g.npartials = 0; g.nranks = 1; g.rank = 0;
  return PCB200_OK;
}

extern "C" int pcb200_timer_start(void) { CK(cudaEventRecord(g.ev_t0, g.st)); return PCB200_OK; }
extern "C" int pcb200_timer_stop(float *ms)
{
  CK(cudaEventRecord(g.ev_t1, g.st));
  CK(cudaEventSynchronize(g.ev_t1));
  CK(cudaEventElapsedTime(ms, g.ev_t0, g.ev_t1));
  return PCB200_OK;
}
extern "C" int pcb200_sync(void) { CK(cudaStreamSynchronize(g.st)); if (g.st_comm) CK(cudaStreamSynchronize(g.st_comm)); return PCB200_OK; }
extern "C" int64_t pcb200_launch_count(void) { const int64_t n = g.launches; g.launches = 0; return n; }

// ------------------------------------------------------------------------
// pcb200_cfl_dt_noalloc: the CFL pre-pass (cfl_prepass -> fixed-order reduction -> dt) for callers that
// must not allocate.  No counterpart in the reference.
// The per-CTA partial maxima, which the regular path keeps in the cudaMalloc'ed g.partials, live in a
// file-scope device array of NOALLOC_MAX_BLOCKS rows of NRED doubles (1 MiB, held by every GPU that loads the library
// until exit).  Every call overwrites the entries it reads (one per CTA of its own launch), so nothing
// has to start at zero; all calls share the array: calls that overlap on the GPU race.
// ------------------------------------------------------------------------
#define NOALLOC_MAX_BLOCKS PCB200_NOALLOC_MAX_BLOCKS   // 512^3 in 32x8x64 blocks
#define NOALLOC_ZC 64
static __device__ __align__(256) double noalloc_scratch[(size_t)NOALLOC_MAX_BLOCKS * NRED];

// second pass: slot RD_DT1 of the nblocks rows in a fixed order, then finish_dt's formula (device code
// may name the array directly; the host side below needs its address from cudaGetSymbolAddress)
__global__ void noalloc_reduce_dt(const double *__restrict__ scratch, int nblocks, double dt1_floor, double *__restrict__ dt_out)
{
  __shared__ double sm[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double v = 0.0;
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x) v = fmax(v, scratch[(size_t)b * NRED + RD_DT1]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) v = fmax(v, sm[w]);
    const double d = fmax(v, dt1_floor);
    dt_out[1] = d; dt_out[0] = 1.0 / d;
  }
}

// device address of the scratch array, looked up once per device with that device current
static int noalloc_scratch_ptr(double **out)
{
  static double *cache[64] = {nullptr};
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(PCB200_EARG, "pcb200_cfl_dt_noalloc: device ordinal out of range");
  if (!cache[dev]) {
    double *p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, noalloc_scratch));
    if (!p) return fail(PCB200_ECUDA, "cudaGetSymbolAddress(noalloc_scratch) returned a null address");
    cache[dev] = p;
  }
  *out = cache[dev];
  return PCB200_OK;
}

extern "C" int pcb200_cfl_dt_noalloc(const double *f_dev, double *dt_out_dev)
{
  if (!g.have_par) return fail(PCB200_ESTATE, "pcb200_cfl_dt_noalloc before pcb200_initialize");
  if (!f_dev || !dt_out_dev) return fail(PCB200_EARG, "pcb200_cfl_dt_noalloc: null device pointer");
  if (((uintptr_t)f_dev & 15) != 0 || ((uintptr_t)dt_out_dev & 7) != 0) return fail(PCB200_EARG, "pcb200_cfl_dt_noalloc: f_dev must be 16-byte aligned, dt_out_dev 8-byte aligned");
  const pcb200_params &P = g.P;
  if (P.lentropy || P.gamma != 1.0) return fail(PCB200_EUNSUPP, "pcb200_cfl_dt_noalloc: the CFL pre-pass covers isothermal runs only (no entropy module, gamma = 1)");
  dim3 blk(TX, TY), grd((P.nx + TX - 1) / TX, (P.ny + TY - 1) / TY, (P.nz + NOALLOC_ZC - 1) / NOALLOC_ZC);
  const long long nb = (long long)grd.x * grd.y * grd.z;
  if (nb > NOALLOC_MAX_BLOCKS)
    return fail(PCB200_EARG, "pcb200_cfl_dt_noalloc: grid needs " + std::to_string(nb) + " blocks of 32x8x64 cells, the static scratch holds " +
                             std::to_string(NOALLOC_MAX_BLOCKS) + " (512^3 cells per rank)");
  CK(cudaSetDevice(g.dev));
  double *scratch = nullptr;
  int rc = noalloc_scratch_ptr(&scratch);
  if (rc) return rc;
  if (P.lmagnetic) cfl_prepass<true><<<grd, blk, 0, g.st>>>(f_dev, scratch, 0, P.nz, NOALLOC_ZC);
  else cfl_prepass<false><<<grd, blk, 0, g.st>>>(f_dev, scratch, 0, P.nz, NOALLOC_ZC);
  const double floor1 = P.dtmax != 0.0 ? 1.0 / P.dtmax : 0.0;
  noalloc_reduce_dt<<<1, 1024, 0, g.st>>>(scratch, (int)nb, floor1, dt_out_dev);
  g.launches += 2;
  CK(cudaGetLastError());
  return PCB200_OK;
}
