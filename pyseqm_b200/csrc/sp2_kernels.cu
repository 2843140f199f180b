// K3b: SP2 density-matrix purification on packed matrices (pack/unpack folded into the first and last
// kernel).  X0 = (hN I - F)/(hN - h1) from Gershgorin bounds, then X <- X^2 or 2X - X^2 by the trace rule
// until two consecutive |tr X - nocc| < eps; result 2 X.  The multiplications run on the FP64 tensor pipe
// (dgemm_kernels.cu); branch choice, trace and convergence flags stay on the device, the host reads one
// counter per iteration (the reference's `while notconverged.any()`).
// reference: seqm/seqm_functions/SP2.py:3-72, pack.py:7-41
#include <dlfcn.h>
#include <cstring>

#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

// ---- NCCL, bound at run time (the library carries no link-time dependency on it; the copy torch already
// loaded is reused).  Only the stable entry points of nccl.h are used; the declarations below restate them.
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
constexpr int NCCL_FLOAT64 = 8;          // ncclDataType_t::ncclFloat64
struct NcclApi {
  int (*GetUniqueId)(NcclUniqueId*);
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int);
  int (*CommDestroy)(NcclComm);
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
  const char* (*GetErrorString)(int);
  bool ok;
};
static NcclApi* nccl_api() {
  static NcclApi api = [] {
    NcclApi a{};
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return a;
    a.GetUniqueId = (int (*)(NcclUniqueId*))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (int (*)(NcclComm))dlsym(h, "ncclCommDestroy");
    a.AllGather = (int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t))dlsym(h, "ncclAllGather");
    a.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.GetErrorString;
    return a;
  }();
  return &api;
}

// what seqm_nccl_comm_create returns: the communicator plus the side stream and events of the pipelined all-gather
constexpr int SP2_MAXSUB = 8;
struct SeqmComm {
  NcclComm comm;
  int rank, nranks;
  cudaStream_t side;
  cudaEvent_t computed[SP2_MAXSUB], gathered, ready;
  // peer-to-peer form: record buffers owned by the library (two step parities), mapped into every rank with CUDA IPC
  bool p2p_tried, p2p;
  size_t rec_cap;                  // doubles per parity
  double* rec_own;                 // cudaMalloc: 2 * rec_cap doubles
  double* rec_peer[16];            // rank q's rec_own as seen from this process (own rank: rec_own)
  double* bar;                     // [2 * nranks] doubles: barrier payload
};


__device__ __forceinline__ int pk2pad(int q, int nh4) { return q < nh4 ? q : nh4 + 4 * (q - nh4); }

__device__ __forceinline__ double blk_reduce(double v, double* red, int op) {   // op 0 sum, 1 min, 2 max
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double u = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? v + u : (op == 1 ? fmin(v, u) : fmax(v, u));
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = red[0];
  for (int q = 1; q < (int)((blockDim.x + 31) >> 5); ++q) s = op == 0 ? s + red[q] : (op == 1 ? fmin(s, red[q]) : fmax(s, red[q]));
  return s;
}

// Initialisation, cut so that ONE large molecule still fills the GPU (config 5 is a single 3000 x 3000 matrix):
//   sp2_bounds_kernel  warp per packed row (8 rows per CTA): Gershgorin interval of the row -> per-CTA min / max
//   sp2_x0_kernel      CTA per chunk of RED_CHUNK elements of X: h1, hN from the per-CTA bounds (every CTA reduces
//                      the same numbers in the same order), X0 = (hN I - F)/(hN - h1), partial trace and |X0|_F^2
//   sp2_step_init_kernel  CTA per molecule: sums the partials in a fixed order, initial state of the fused loop
constexpr int SP2_ROWS = 8;

// The reference purifies pack(F[notconverged]) (scf_loop.py + pack.py:28): every molecule of the call zero padded to the
// LARGEST ACTIVE molecule, and SP2.py:30-41 takes its Gershgorin bounds and traces over the padded matrix.  A zero row is
// an orbital of energy 0 that never couples: all npad = padm - m of them carry the same scalar xp = hN/(hN - h1), which
// follows the branch decisions (xp^2 or 2 xp - xp^2) and enters tr X and tr X^2 with weight npad.  Tracking that scalar
// reproduces the reference's bounds, branches and stopping step exactly without touching the matrices.
__global__ void sp2_accumulate_kernel(const int* __restrict__ live, int32_t* __restrict__ total) { *total += *live; }

__global__ void sp2_padmax_kernel(int nmol, const int32_t* __restrict__ nheavy, const int32_t* __restrict__ nhydro,
                                  const uint8_t* __restrict__ active, int* __restrict__ padm) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nmol || (active && !active[b])) return;
  atomicMax(padm, 4 * nheavy[b] + nhydro[b]);
}

__global__ void __launch_bounds__(256)
sp2_bounds_kernel(int S, int rchunks, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                  const int32_t* __restrict__ nhydro, const uint8_t* __restrict__ active, double* __restrict__ bounds,
                  int32_t* __restrict__ dims, const int* __restrict__ padm) {
  const int mol = blockIdx.x / rchunks, chunk = blockIdx.x - mol * rchunks;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (chunk == 0 && threadIdx.x == 0) dims[mol] = m;
  if ((active && !active[mol]) || m <= 0) return;
  const int n4 = 4 * S;
  const double* Fm = F + (size_t)mol * n4 * n4;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = chunk * SP2_ROWS + warp;
  double lo = 1.0e300, hi = -1.0e300;
  if (i < m) {
    const size_t row = (size_t)pk2pad(i, nh4) * n4;
    double rad = 0.0, aii = 0.0;
    for (int j = lane; j < m; j += 32) {
      const double a = Fm[row + pk2pad(j, nh4)];
      if (j == i) aii = a; else rad += fabs(a);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      rad += __shfl_xor_sync(0xffffffffu, rad, o);
      aii += __shfl_xor_sync(0xffffffffu, aii, o);
    }
    lo = aii - rad;
    hi = aii + rad;
  }
  __shared__ double slo[SP2_ROWS], shi[SP2_ROWS];
  if (lane == 0) { slo[warp] = lo; shi[warp] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < SP2_ROWS; ++q) { lo = fmin(lo, slo[q]); hi = fmax(hi, shi[q]); }
    if (chunk == 0 && m < *padm) { lo = fmin(lo, 0.0); hi = fmax(hi, 0.0); }      // the zero rows of the reference's padding
    bounds[((size_t)mol * rchunks + chunk) * 2] = lo;
    bounds[((size_t)mol * rchunks + chunk) * 2 + 1] = hi;
  }
}

__global__ void __launch_bounds__(256)
sp2_x0_kernel(int S, int ld, int rchunks, int chunks, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
              const int32_t* __restrict__ nhydro, const uint8_t* __restrict__ active, const double* __restrict__ bounds,
              double* __restrict__ X, double* __restrict__ part, const int* __restrict__ padm, double* __restrict__ padx) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if ((active && !active[mol]) || m <= 0) return;
  __shared__ double red[8];
  const int mr = (m + SP2_ROWS - 1) / SP2_ROWS;       // row chunks that hold rows of this molecule
  double lo = 1.0e300, hi = -1.0e300;
  for (int q = threadIdx.x; q < mr; q += blockDim.x) {
    lo = fmin(lo, bounds[((size_t)mol * rchunks + q) * 2]);
    hi = fmax(hi, bounds[((size_t)mol * rchunks + q) * 2 + 1]);
  }
  const double h1 = blk_reduce(lo, red, 1);
  const double hN = blk_reduce(hi, red, 2);
  const double sc = hN - h1;
  const int n4 = 4 * S;
  const double* Fm = F + (size_t)mol * n4 * n4;
  double* Xm = X + (size_t)mol * ld * ld;
  const size_t tot = (size_t)ld * ld;
  const size_t e0 = (size_t)chunk * RED_CHUNK, e1 = e0 + RED_CHUNK < tot ? e0 + RED_CHUNK : tot;
  double tr = 0.0, fr = 0.0;
  for (size_t idx = e0 + threadIdx.x; idx < e1; idx += blockDim.x) {
    int i, j;
    row_col(idx, ld, i, j);
    double x = 0.0;
    if (i < m && j < m) {
      const double a = Fm[(size_t)pk2pad(i, nh4) * n4 + pk2pad(j, nh4)];
      x = ((i == j ? hN : 0.0) - a) / sc;
      if (i == j) tr += x;
      fr += x * x;
    }
    Xm[idx] = x;
  }
  tr = blk_reduce(tr, red, 0);
  fr = blk_reduce(fr, red, 0);
  if (threadIdx.x == 0) {
    if (chunk == 0) {
      const int npad = *padm > m ? *padm - m : 0;
      const double xp = hN / sc;
      padx[mol] = xp;
      tr += npad * xp;
      fr += npad * xp * xp;
    }
    part[((size_t)mol * chunks + chunk) * 2] = tr;
    part[((size_t)mol * chunks + chunk) * 2 + 1] = fr;
  }
}

// ---------------------------------------------------------------------------------------------
// Fused SP2 step: one launch per purification step.
//   prologue   every CTA of molecule b sums the (fixed-order) partial traces the previous step left behind:
//              t1 = tr X, t2 = |X|_F^2 = tr X^2 (X symmetric), takes the branch of the trace rule and the
//              convergence decision -- all CTAs of a molecule compute the same scalars from the same numbers
//   main loop  Y = X X for one 64x64 tile of the UPPER block triangle (DMMA m8n8k4, cp.async double buffering)
//   epilogue   Xnew = Y or 2X - Y, staged through shared memory and written to the other buffer as the tile and
//              its transpose (coalesced both ways); partial sums of diag(Xnew) and |Xnew|_F^2 for the next step
// Per-molecule scalars are double buffered by step parity (tile 0 writes, everybody reads the other copy).
// ---------------------------------------------------------------------------------------------
constexpr int SBM = 64, SBK = 16;
constexpr int SLDA = SBK + 4, SLDB = SBM + 4, SLDT = SBM + 1;

struct Sp2Step {       // per molecule and parity
  double e0, e1;       // |tr X - nocc| after the last two steps
  double xp;           // value of the decoupled padding orbitals (reference zero padding, see sp2_padmax_kernel)
  int live, niter, buf, pad;      // pad = number of padding orbitals
};

__device__ __forceinline__ void s_cp_async16(void* smem, const void* gmem, bool pred) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  const int bytes = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(sa), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void s_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// SH = true is the tile-sharded form (single large system over several GPUs, seqm_sp2_density_nccl): this launch
// covers the tiles [tile0, tile0 + gridDim.x) of the enumeration, the new tile and its two partial sums go into
// a contiguous 4098-double RECORD (rec + (b * gridDim.x + blockIdx.x) * 4098) instead of X; after the all-gather
// of the records sp2_unpack_kernel writes every tile, its transpose and the partial sums where the SH = false
// form puts them directly.  `lead` marks the launch whose block 0 owns the per-molecule scalars.
constexpr int SP2_REC = SBM * SBM + 2;
constexpr int SP2_MAXPEER = 16;
struct Sp2Peers { double* p[SP2_MAXPEER]; int n; };

template <bool SH>
__global__ void __launch_bounds__(256)
sp2_step_kernel(int nmol, int b0, int ld, int T, int step, double eps, const int32_t* __restrict__ dims,
                const double* __restrict__ occ, double* __restrict__ Xbuf0, double* __restrict__ Xbuf1,
                long long stride, Sp2Step* __restrict__ state, double* __restrict__ partial, int* __restrict__ nlive,
                int tile0, int lead, Sp2Peers peers, size_t recoff) {
  const int b = b0 + blockIdx.y;
  const Sp2Step sin = state[(size_t)(step & 1) * nmol + b];
  const int ntile = T * (T + 1) / 2;
  const int tileid = SH ? tile0 + (int)blockIdx.x : (int)blockIdx.x;
  const bool leader = SH ? (lead && blockIdx.x == 0) : (blockIdx.x == 0);
  // the record of this tile inside every destination buffer (own GPU and, in the peer-to-peer form, every peer)
  const size_t myoff = SH ? recoff + ((size_t)b * gridDim.x + blockIdx.x) * SP2_REC : 0;
  if (!sin.live) {      // keep both parities of a finished molecule identical
    if (leader && threadIdx.x == 0) state[(size_t)((step + 1) & 1) * nmol + b] = sin;
    return;
  }
  const int nn = dims[b];
  const int TT = (nn + SBM - 1) / SBM;                      // tile rows of this molecule
  // tile id -> (ti <= tj) in the batch-wide enumeration of the upper block triangle of a T x T grid
  int ti = 0, rem = tileid;
  if (SH && tileid >= ntile) { ti = T; rem = 0; }           // padding block of the last rank
  while (ti < T && rem >= T - ti) { rem -= T - ti; ++ti; }
  const int tj = ti + rem;
  const double* pin = partial + ((size_t)(step & 1) * nmol + b) * 2 * ntile;
  double* pout = partial + ((size_t)((step + 1) & 1) * nmol + b) * 2 * ntile;
  double t1 = 0.0, t2 = 0.0;
  for (int q = 0; q < ntile; ++q) { t1 += pin[2 * q]; t2 += pin[2 * q + 1]; }
  if (step > 0) { t1 += sin.pad * sin.xp; t2 += sin.pad * sin.xp * sin.xp; }      // step 0: already in the initial sums
  const double oc = occ[b];
  const bool square = fabs(t2 - oc) < fabs(2.0 * t1 - t2 - oc);
  const double tn = square ? t2 : 2.0 * t1 - t2;
  const int tid = threadIdx.x;
  if (leader && tid == 0) {
    Sp2Step so;
    so.e1 = sin.e0;
    so.e0 = fabs(tn - oc);
    so.niter = sin.niter + 1;
    so.buf = (sin.buf ^ 1);
    so.live = ((so.e0 < eps) && (so.e1 < eps)) ? 0 : 1;
    so.pad = sin.pad;
    so.xp = square ? sin.xp * sin.xp : 2.0 * sin.xp - sin.xp * sin.xp;
    state[(size_t)((step + 1) & 1) * nmol + b] = so;
    if (so.live) atomicAdd(nlive, 1);
  }
  if (ti >= TT || tj >= TT) {           // tile outside this (smaller) molecule: contributes nothing
    if (tid == 0) {
      if (SH) {
        for (int q = 0; q < peers.n; ++q) { peers.p[q][myoff + SBM * SBM] = 0.0; peers.p[q][myoff + SBM * SBM + 1] = 0.0; }
      }
      else { pout[2 * blockIdx.x] = 0.0; pout[2 * blockIdx.x + 1] = 0.0; }
    }
    return;
  }
  const double* Xin = (sin.buf ? Xbuf1 : Xbuf0) + (size_t)b * stride;
  double* Xout = (sin.buf ? Xbuf0 : Xbuf1) + (size_t)b * stride;
  const int m0 = ti * SBM, n0 = tj * SBM;
  __shared__ __align__(16) double sm[2 * SBM * SLDA + 2 * SBK * SLDB];
  double* As0 = sm;
  double* Bs0 = sm + 2 * SBM * SLDA;
  const int lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 1) * 16, wn = (warp & 1) * 32;
  const int g = lane >> 2, t = lane & 3;
  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  auto load_tiles = [&](int buf, int k0) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256;
      const int r = idx >> 3, c2 = (idx & 7) * 2;
      const int gr = m0 + r, gc = k0 + c2;
      const bool ok = gr < nn && gc < nn;
      s_cp_async16(&As0[buf * SBM * SLDA + r * SLDA + c2], Xin + (size_t)(ok ? gr : 0) * ld + (ok ? gc : 0), ok);
    }
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256;
      const int r = idx >> 5, c2 = (idx & 31) * 2;
      const int gr = k0 + r, gc = n0 + c2;
      const bool ok = gr < nn && gc < nn;
      s_cp_async16(&Bs0[buf * SBK * SLDB + r * SLDB + c2], Xin + (size_t)(ok ? gr : 0) * ld + (ok ? gc : 0), ok);
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
  // diagonal tiles: the two 16 x 32 warp tiles strictly below the diagonal are not multiplied - the epilogue mirrors
  // the upper half (bit-identical: X is exactly symmetric, so both halves are the same sums in the same order)
  const bool below = (ti == tj) && (wm >= wn + 32);
  const int nk = (nn + SBK - 1) / SBK;
  load_tiles(0, 0);
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) {
      load_tiles(buf ^ 1, (kt + 1) * SBK);
      asm volatile("cp.async.wait_group 1;\n" ::);
    } else {
      asm volatile("cp.async.wait_group 0;\n" ::);
    }
    __syncthreads();
    const double* as = As0 + buf * SBM * SLDA;
    const double* bs = Bs0 + buf * SBK * SLDB;
    if (!below) {
#pragma unroll
      for (int kk = 0; kk < SBK; kk += 4) {
        double af[2], bf[4];
#pragma unroll
        for (int i = 0; i < 2; ++i) af[i] = as[(wm + i * 8 + g) * SLDA + kk + t];
#pragma unroll
        for (int j = 0; j < 4; ++j) bf[j] = bs[(kk + t) * SLDB + wn + j * 8 + g];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) s_dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
      }
    }
    __syncthreads();
  }
  // ---- epilogue: Xnew tile into shared memory (the staging buffers are free now)
  double* Ts = sm;                      // [64][65]
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int rl = wm + i * 8 + g;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cl = wn + j * 8 + 2 * t;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = m0 + rl, c = n0 + cl + h;
        double y = acc[i][j][h];
        if (r < nn && c < nn) {
          if (!square) y = 2.0 * Xin[(size_t)r * ld + c] - y;
        } else {
          y = 0.0;
        }
        Ts[rl * SLDT + cl + h] = y;
      }
    }
  }
  __syncthreads();
  double ptr = 0.0, pfr = 0.0;
  for (int idx = tid; idx < SBM * SBM; idx += 256) {
    const int rl = idx >> 6, cl = idx & 63;
    const int r = m0 + rl, c = n0 + cl;
    const bool mir = (ti == tj) && (rl > cl);
    if (SH) {
      const double y = mir ? Ts[cl * SLDT + rl] : Ts[rl * SLDT + cl];          // zero outside the matrix
      for (int q = 0; q < peers.n; ++q) peers.p[q][myoff + idx] = y;      // NVLink stores for q != own rank
      pfr += y * y;
      if (ti == tj && rl == cl) ptr += y;
      continue;
    }
    if (r < nn && c < nn) {
      const double y = mir ? Ts[cl * SLDT + rl] : Ts[rl * SLDT + cl];
      Xout[(size_t)r * ld + c] = y;
      pfr += y * y;
      if (ti == tj && rl == cl) ptr += y;
    }
    if (ti != tj) {                     // transposed tile: row n0 + rl, column m0 + cl
      const int r2 = n0 + rl, c2 = m0 + cl;
      if (r2 < nn && c2 < nn) Xout[(size_t)r2 * ld + c2] = Ts[cl * SLDT + rl];
    }
  }
  if (ti != tj) pfr *= 2.0;
  __shared__ double red[8];
  ptr = blk_reduce(ptr, red, 0);
  pfr = blk_reduce(pfr, red, 0);
  if (tid == 0) {
    if (SH) {
      for (int q = 0; q < peers.n; ++q) { peers.p[q][myoff + SBM * SBM] = ptr; peers.p[q][myoff + SBM * SBM + 1] = pfr; }
    }
    else { pout[2 * blockIdx.x] = ptr; pout[2 * blockIdx.x + 1] = pfr; }
  }
}

// Sharded form, after the all-gather: records of the tiles [tile0, tile0 + gridDim.x) of every molecule ->
// X (tile and transposed tile) of the buffer this step writes, and the partial sums of the next step.
__global__ void __launch_bounds__(256)
sp2_unpack_kernel(int nmol, int ld, int T, int step, const int32_t* __restrict__ dims, double* __restrict__ Xbuf0,
                  double* __restrict__ Xbuf1, long long stride, const Sp2Step* __restrict__ state,
                  double* __restrict__ partial, int tile0, const double* __restrict__ rec, int tpr) {
  // tpr > 0: one launch over the segments of ALL ranks, layout [rank][molecule][tpr] (peer-to-peer form)
  const int b = blockIdx.y;
  const Sp2Step sin = state[(size_t)(step & 1) * nmol + b];
  if (!sin.live) return;
  const int ntile = T * (T + 1) / 2;
  const int tileid = tile0 + (int)blockIdx.x;
  if (tileid >= ntile) return;
  const double* myrec = tpr > 0
      ? rec + (((size_t)(blockIdx.x / tpr) * gridDim.y + b) * tpr + (blockIdx.x % tpr)) * SP2_REC
      : rec + ((size_t)b * gridDim.x + blockIdx.x) * SP2_REC;
  double* pout = partial + ((size_t)((step + 1) & 1) * nmol + b) * 2 * ntile;
  const int tid = threadIdx.x;
  if (tid == 0) { pout[2 * tileid] = myrec[SBM * SBM]; pout[2 * tileid + 1] = myrec[SBM * SBM + 1]; }
  const int nn = dims[b];
  const int TT = (nn + SBM - 1) / SBM;
  int ti = 0, rem = tileid;
  while (ti < T && rem >= T - ti) { rem -= T - ti; ++ti; }
  const int tj = ti + rem;
  if (ti >= TT || tj >= TT) return;
  double* Xout = (sin.buf ? Xbuf0 : Xbuf1) + (size_t)b * stride;
  const int m0 = ti * SBM, n0 = tj * SBM;
  __shared__ double Ts[SBM * SLDT];
  for (int idx = tid; idx < SBM * SBM; idx += 256) Ts[(idx >> 6) * SLDT + (idx & 63)] = myrec[idx];
  __syncthreads();
  for (int idx = tid; idx < SBM * SBM; idx += 256) {
    const int rl = idx >> 6, cl = idx & 63;
    const int r = m0 + rl, c = n0 + cl;
    if (r < nn && c < nn) Xout[(size_t)r * ld + c] = Ts[rl * SLDT + cl];
    if (ti != tj) {
      const int r2 = n0 + rl, c2 = m0 + cl;
      if (r2 < nn && c2 < nn) Xout[(size_t)r2 * ld + c2] = Ts[cl * SLDT + rl];
    }
  }
}

// initial per-molecule scalars of the fused loop from the partial sums sp2_x0_kernel left (X0 in buffer 0)
__global__ void __launch_bounds__(256)
sp2_step_init_kernel(int nmol, int ntile, int chunks, const int32_t* __restrict__ dims,
                     const int32_t* __restrict__ nocc, const uint8_t* __restrict__ active,
                     const double* __restrict__ part, Sp2Step* __restrict__ state, double* __restrict__ partial,
                     double* __restrict__ occ, const int* __restrict__ padm, const double* __restrict__ padx) {
  const int b = blockIdx.x;
  __shared__ double red[8];
  const bool live = (!active || active[b]) && dims[b] > 0;
  double tr = 0.0, fr = 0.0;
  if (live) {
    for (int q = threadIdx.x; q < chunks; q += blockDim.x) {
      tr += part[((size_t)b * chunks + q) * 2];
      fr += part[((size_t)b * chunks + q) * 2 + 1];
    }
  }
  tr = blk_reduce(tr, red, 0);
  fr = blk_reduce(fr, red, 0);
  for (int q = threadIdx.x; q < 2 * ntile; q += blockDim.x) partial[(size_t)b * 2 * ntile + q] = 0.0;
  __syncthreads();
  if (threadIdx.x == 0) {
    const double oc = (double)nocc[b];
    Sp2Step s;
    s.e0 = fabs(tr - oc); s.e1 = s.e0; s.live = live ? 1 : 0; s.niter = 0; s.buf = 0;
    s.pad = (live && *padm > dims[b]) ? *padm - dims[b] : 0;
    s.xp = s.pad ? padx[b] : 0.0;
    state[b] = s;
    s.live = 0;
    state[(size_t)nmol + b] = s;
    occ[b] = oc;
    partial[(size_t)b * 2 * ntile] = tr;
    partial[(size_t)b * 2 * ntile + 1] = fr;
  }
}

// D = 2 X from the buffer each molecule finished in
__global__ void __launch_bounds__(256)
sp2_finish2_kernel(int S, int ld, const double* __restrict__ X0, const double* __restrict__ X1, long long stride,
                   const int32_t* __restrict__ nheavy, const int32_t* __restrict__ dims,
                   const uint8_t* __restrict__ active, const Sp2Step* __restrict__ sa, const Sp2Step* __restrict__ sb,
                   int b0, double* __restrict__ D, int32_t* __restrict__ niter) {
  const int mol = b0 + blockIdx.y;
  if (active && !active[mol]) return;
  // the entry with the larger iteration count is the last one written for this molecule
  const Sp2Step s = sa[mol].niter >= sb[mol].niter ? sa[mol] : sb[mol];
  const int m = dims[mol], nh4 = 4 * nheavy[mol], n4 = 4 * S;
  if (niter && blockIdx.x == 0 && threadIdx.x == 0) niter[mol] = s.niter;
  const double* Xm = (s.buf ? X1 : X0) + (size_t)mol * stride;
  double* Dm = D + (size_t)mol * n4 * n4;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < m * m; idx += gridDim.x * blockDim.x) {
    const int i = idx / m, j = idx - i * m;
    Dm[(size_t)pk2pad(i, nh4) * n4 + pk2pad(j, nh4)] = 2.0 * Xm[(size_t)i * ld + j];
  }
}

void* g_scf_sp2_comm = nullptr;      // set by seqm_scf_sp2_comm(): the SCF driver's SP2 runs tile-sharded
int g_scf_sp2_nranks = 0;

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static int sp2_ld(int mmax) { return (mmax + 3) & ~3; }
static int sp2_rchunks(int ld) { return (ld + SP2_ROWS - 1) / SP2_ROWS; }
static int sp2_tiles(int ld) { const int T = (ld + SBM - 1) / SBM; return T * (T + 1) / 2; }

}  // namespace seqm

using namespace seqm;

extern "C" {

// tiles per rank and sub-chunk: the batch-wide enumeration [0, ntile) is cut into nranks contiguous ranges of
// tpr = nsub * tps tiles (the last ranks may own padding tiles >= ntile)
void seqm_sp2_shard_plan(int ntile, int nranks, int nsub, int rank, int32_t* out4) {
  if (nranks < 1) nranks = 1;
  if (nsub < 1) nsub = 1;
  int tpr = (ntile + nranks - 1) / nranks;
  const int tps = (tpr + nsub - 1) / nsub;
  tpr = tps * nsub;
  int lo = rank * tpr, hi = lo + tpr;
  if (lo > ntile) lo = ntile;
  if (hi > ntile) hi = ntile;
  out4[0] = tpr; out4[1] = tps; out4[2] = lo; out4[3] = hi;      // owned real tiles: [lo, hi)
}

static int g_sp2_nsub = 0;      // 0: automatic
void seqm_sp2_nccl_subchunks(int n) { g_sp2_nsub = n < 0 ? 0 : (n > SP2_MAXSUB ? SP2_MAXSUB : n); }
// automatic choice: a piece must still fill the GPU (148 SMs x 4 resident CTAs of sp2_step_kernel), otherwise
// the multiplication loses more to idle SMs than the overlap of the exchange wins
static int sp2_auto_nsub(int ntile, int nranks, int nmol) {
  const long long per_rank = ((long long)(ntile + nranks - 1) / nranks) * nmol;
  long long n = per_rank / (148 * 4);
  return n < 1 ? 1 : (n > 4 ? 4 : (int)n);
}

static size_t sp2_core_bytes(int64_t nmol, int64_t molsize) {
  const size_t ld = (size_t)4 * molsize;       // upper bound of the packed size
  const size_t ntile = (size_t)sp2_tiles((int)ld);
  return 2 * al256((size_t)nmol * ld * ld * 8) + al256((size_t)nmol * sp2_rchunks((int)ld) * 16) + al256(nmol * 4) +
         al256(2 * nmol * sizeof(Sp2Step)) + al256(2 * nmol * 2 * ntile * 8) + 2 * al256(nmol * 8) + 256 + 256 +
         al256((size_t)nmol * red_chunks(ld * ld) * 16);
}
static size_t sp2_record_bytes(int64_t nmol, int64_t molsize, int nranks) {
  const size_t ntile = (size_t)sp2_tiles(4 * (int)molsize);
  const size_t tpr_bound = (ntile + nranks - 1) / nranks + SP2_MAXSUB;      // >= tiles per rank for any sub-chunk count
  return al256((size_t)nranks * tpr_bound * nmol * SP2_REC * 8);
}

size_t seqm_sp2_workspace(int64_t nmol, int64_t molsize) { return sp2_core_bytes(nmol, molsize); }
size_t seqm_sp2_nccl_workspace(int64_t nmol, int64_t molsize, int nranks) {
  return sp2_core_bytes(nmol, molsize) + sp2_record_bytes(nmol, molsize, nranks < 1 ? 1 : nranks);
}

int seqm_nccl_unique_id(void* id128) {
  SEQM_REQUIRE(id128, "seqm_nccl_unique_id: null pointer");
  NcclApi* a = nccl_api();
  if (!a->ok) return fail(SEQM_ECUDA, "seqm_nccl_unique_id: libnccl.so.2 not found");
  NcclUniqueId id;
  const int r = a->GetUniqueId(&id);
  if (r) return fail(SEQM_ECUDA, a->GetErrorString(r));
  memcpy(id128, &id, 128);
  return SEQM_OK;
}

int seqm_nccl_comm_create(const void* id128, int rank, int nranks, void** comm_out) {
  SEQM_REQUIRE(id128 && comm_out && nranks >= 1 && rank >= 0 && rank < nranks, "seqm_nccl_comm_create: bad argument");
  NcclApi* a = nccl_api();
  if (!a->ok) return fail(SEQM_ECUDA, "seqm_nccl_comm_create: libnccl.so.2 not found");
  NcclUniqueId id;
  memcpy(&id, id128, 128);
  SeqmComm* c = new SeqmComm();
  c->rank = rank; c->nranks = nranks;
  c->p2p_tried = c->p2p = false; c->rec_cap = 0; c->rec_own = nullptr; c->bar = nullptr;
  for (int i = 0; i < 16; ++i) c->rec_peer[i] = nullptr;
  const int r = a->CommInitRank(&c->comm, nranks, id, rank);
  if (r) { delete c; return fail(SEQM_ECUDA, a->GetErrorString(r)); }
  SEQM_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
  for (int i = 0; i < SP2_MAXSUB; ++i) SEQM_CUDA(cudaEventCreateWithFlags(&c->computed[i], cudaEventDisableTiming));
  SEQM_CUDA(cudaEventCreateWithFlags(&c->gathered, cudaEventDisableTiming));
  SEQM_CUDA(cudaEventCreateWithFlags(&c->ready, cudaEventDisableTiming));
  *comm_out = c;
  return SEQM_OK;
}

int seqm_nccl_comm_destroy(void* comm) {
  if (!comm) return SEQM_OK;
  SeqmComm* c = (SeqmComm*)comm;
  if (g_scf_sp2_comm == comm) { g_scf_sp2_comm = nullptr; g_scf_sp2_nranks = 0; }
  cudaDeviceSynchronize();
  if (c->rec_own) {
    for (int q = 0; q < c->nranks && q < 16; ++q)
      if (q != c->rank && c->rec_peer[q]) cudaIpcCloseMemHandle(c->rec_peer[q]);
    cudaFree(c->rec_own);
  }
  if (c->bar) cudaFree(c->bar);
  nccl_api()->CommDestroy(c->comm);
  cudaStreamDestroy(c->side);
  for (int i = 0; i < SP2_MAXSUB; ++i) cudaEventDestroy(c->computed[i]);
  cudaEventDestroy(c->gathered);
  cudaEventDestroy(c->ready);
  delete c;
  return SEQM_OK;
}

// comm: handle of seqm_nccl_comm_create, or NULL with emulate_ranks >= 1: the tile ranges of `emulate_ranks`
// ranks are computed one after the other on this GPU through the same records (test form of the sharded path)
void seqm_scf_sp2_comm(void* comm, int emulate_ranks) {
  g_scf_sp2_comm = comm;
  g_scf_sp2_nranks = comm ? ((SeqmComm*)comm)->nranks : emulate_ranks;
}

static int g_sp2_p2p = 1;
void seqm_sp2_nccl_p2p(int on) { g_sp2_p2p = on; }
static size_t seg_full(int B, int tpr) { return (size_t)B * tpr * SP2_REC; }

// Record buffers of the peer-to-peer form: allocated by the library (cudaMalloc, two step parities), their CUDA IPC
// handles all-gathered through NCCL and opened by every rank.  Collective: every rank calls it with the same size.
// Returns false (and the caller uses ncclAllGather on the torch-owned records) if any rank could not map a peer.
static bool sp2_p2p_ready(SeqmComm* cm, size_t doubles) {
  if (!g_sp2_p2p || cm->nranks > SP2_MAXPEER) return false;
  if (cm->p2p_tried && (!cm->p2p || cm->rec_cap >= doubles)) return cm->p2p;
  NcclApi* api = nccl_api();
  const int n = cm->nranks;
  if (cm->rec_own) {      // grow: unmap, free, start again (all ranks take this branch together)
    cudaDeviceSynchronize();
    for (int q = 0; q < n; ++q)
      if (q != cm->rank && cm->rec_peer[q]) cudaIpcCloseMemHandle(cm->rec_peer[q]);
    if (cm->bar) {      // no rank frees its buffer before every peer has unmapped it
      double* cb = cm->bar + (size_t)13 * n;
      if (api->AllGather(cb + cm->rank, cb, 1, NCCL_FLOAT64, cm->comm, cm->side) == 0) cudaStreamSynchronize(cm->side);
    }
    cudaFree(cm->rec_own);
    cm->rec_own = nullptr;
  }
  cm->p2p_tried = true;
  cm->p2p = false;
  cm->rec_cap = (doubles + 511) & ~(size_t)511;
  bool ok = cudaMalloc(&cm->rec_own, 2 * cm->rec_cap * 8) == cudaSuccess;
  if (!cm->bar) ok = ok && cudaMalloc(&cm->bar, (size_t)(2 * n + 16 * n + 2 * n) * 8) == cudaSuccess;
  // exchange {ok flag, 64-byte IPC handle} = 9 doubles per rank, staged through cm->bar + 2n
  double* xbuf = cm->bar ? cm->bar + 2 * n : nullptr;
  double hostx[SP2_MAXPEER * 9];
  memset(hostx, 0, sizeof(hostx));
  cudaIpcMemHandle_t h;
  memset(&h, 0, sizeof(h));
  if (ok) ok = cudaIpcGetMemHandle(&h, cm->rec_own) == cudaSuccess;
  double mine[9];
  mine[0] = ok ? 1.0 : 0.0;
  memcpy(mine + 1, &h, 64);
  if (!xbuf) { cudaGetLastError(); return false; }
  cudaMemcpy(xbuf + 9 * cm->rank, mine, 72, cudaMemcpyHostToDevice);
  if (api->AllGather(xbuf + 9 * cm->rank, xbuf, 9, NCCL_FLOAT64, cm->comm, cm->side)) { cudaGetLastError(); return false; }
  cudaStreamSynchronize(cm->side);
  cudaMemcpy(hostx, xbuf, (size_t)n * 72, cudaMemcpyDeviceToHost);
  bool all = true;
  for (int q = 0; q < n; ++q) all = all && hostx[9 * q] == 1.0;
  double opened = 1.0;
  for (int q = 0; q < n && all; ++q) {
    if (q == cm->rank) { cm->rec_peer[q] = cm->rec_own; continue; }
    cudaIpcMemHandle_t hq;
    memcpy(&hq, hostx + 9 * q + 1, 64);
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, hq, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { opened = 0.0; cm->rec_peer[q] = nullptr; }
    else cm->rec_peer[q] = (double*)ptr;
  }
  // second round: did every rank map every peer?
  double* flags = xbuf + 9 * n;
  cudaMemcpy(flags + cm->rank, &opened, 8, cudaMemcpyHostToDevice);
  if (api->AllGather(flags + cm->rank, flags, 1, NCCL_FLOAT64, cm->comm, cm->side)) { cudaGetLastError(); return false; }
  cudaStreamSynchronize(cm->side);
  double hf[SP2_MAXPEER];
  cudaMemcpy(hf, flags, (size_t)n * 8, cudaMemcpyDeviceToHost);
  for (int q = 0; q < n; ++q) all = all && hf[q] == 1.0;
  cudaGetLastError();      // a failed IPC call must not poison later launch checks
  cm->p2p = all;
  return all;
}

// shared implementation; nranks == 0: the single-GPU form (tiles written straight into X)
// async_steps > 0: exactly that many (rounded up to even) purification steps are enqueued and nothing is read back -
// no host synchronisation, no host allocation, capturable in a CUDA graph; d_unfinished (device) receives the number of
// molecules still live after the last step.  async_steps == 0: the host reads the per-step live counters back; the
// first read comes after as many steps as the previous call needed (g_sp2_hint), so a steady-state call (MD, SCF)
// synchronises once.
static std::atomic<int> g_sp2_hint{2};
static std::atomic<unsigned long long> g_sp2_steps{0};      // purification steps enqueued since load (seqm_sp2_step_count)
constexpr int SP2_NCOUNT = 64;      // per-step live counters kept on the device (256 bytes of the workspace)
static int sp2_run(SeqmComm* cm, int nranks, int64_t nmol, int64_t molsize, int mmax, const double* F,
                   const int32_t* nheavy, const int32_t* nhydro, const int32_t* nocc, double eps,
                   const uint8_t* active, double* D, int32_t* niter, void* work, void* stream,
                   int async_steps = 0, int32_t* d_unfinished = nullptr) {
  SEQM_REQUIRE(F && nheavy && nhydro && nocc && D && work && nmol >= 0 && molsize > 0 && molsize < 16384, "seqm_sp2_density: bad argument");
  if (nmol == 0) return SEQM_OK;
  SEQM_REQUIRE(mmax >= 1 && mmax <= 4 * molsize, "seqm_sp2_density: mmax");
  const bool sharded = nranks >= 1;
  SEQM_REQUIRE(!sharded || nmol <= 65535, "seqm_sp2_density_nccl: at most 65535 molecules");
  cudaStream_t st = (cudaStream_t)stream;
  if (eps < 1.0e-7) eps = 1.0e-7;      // SP2.py:21-24 (fp64 branch)
  if (eps > 1.0e-3) eps = 1.0e-3;
  const int B = (int)nmol, S = (int)molsize, ld = sp2_ld(mmax);
  const size_t ldmax = (size_t)4 * molsize;
  const int T = (ld + SBM - 1) / SBM, ntile = T * (T + 1) / 2;
  char* base = (char*)work;
  double* X0 = (double*)base; base += al256((size_t)nmol * ldmax * ldmax * 8);
  double* X1 = (double*)base; base += al256((size_t)nmol * ldmax * ldmax * 8);
  double* bounds = (double*)base; base += al256((size_t)nmol * sp2_rchunks((int)ldmax) * 16);
  int32_t* dims = (int32_t*)base; base += al256(nmol * 4);
  Sp2Step* state = (Sp2Step*)base; base += al256(2 * nmol * sizeof(Sp2Step));
  double* partial = (double*)base; base += al256((size_t)2 * nmol * 2 * sp2_tiles((int)ldmax) * 8);
  double* occ = (double*)base; base += al256(nmol * 8);
  double* padx = (double*)base; base += al256(nmol * 8);
  int* padm = (int*)base; base += 256;
  int* nlive = (int*)base; base += 256;      // [SP2_NCOUNT] per-step live counters
  double* part0 = (double*)base; base += al256((size_t)nmol * red_chunks(ldmax * ldmax) * 16);
  double* records = (double*)base;           // sharded form only: [sub][rank][molecule][tps] records
  const long long stride = (long long)ld * ld;
  int32_t plan[4] = {0, 0, 0, 0};
  const bool want_p2p = cm && nranks > 1 && nranks <= SP2_MAXPEER && g_sp2_p2p;
  const int nsub = (!sharded || want_p2p) ? 1 : (g_sp2_nsub > 0 ? g_sp2_nsub : sp2_auto_nsub(ntile, nranks, B));
  const int myrank = cm ? cm->rank : 0;
  if (sharded) seqm_sp2_shard_plan(ntile, nranks, nsub, myrank, plan);
  const int tpr = plan[0], tps = plan[1];
  const size_t seg = (size_t)B * tps * SP2_REC;      // doubles of one (sub, rank) segment
  NcclApi* api = cm ? nccl_api() : nullptr;
  prof_mark("sp2_init", st);
  const int rchunks = sp2_rchunks(ld), xchunks = red_chunks((size_t)ld * ld);
  SEQM_REQUIRE((long long)B * rchunks < (1LL << 31) && (long long)B * xchunks < (1LL << 31), "seqm_sp2_density: batch too large");
  SEQM_CUDA(cudaMemsetAsync(padm, 0, sizeof(int), st));
  sp2_padmax_kernel<<<cdiv(B, 256), 256, 0, st>>>(B, nheavy, nhydro, active, padm);
  SEQM_CHECK_LAUNCH("sp2_padmax_kernel");
  sp2_bounds_kernel<<<(unsigned)((size_t)B * rchunks), 256, 0, st>>>(S, rchunks, F, nheavy, nhydro, active, bounds, dims, padm);
  SEQM_CHECK_LAUNCH("sp2_bounds_kernel");
  sp2_x0_kernel<<<(unsigned)((size_t)B * xchunks), 256, 0, st>>>(S, ld, rchunks, xchunks, F, nheavy, nhydro, active, bounds, X0, part0, padm, padx);
  SEQM_CHECK_LAUNCH("sp2_x0_kernel");
  sp2_step_init_kernel<<<B, 256, 0, st>>>(B, ntile, xchunks, dims, nocc, active, part0, state, partial, occ, padm, padx);
  SEQM_CHECK_LAUNCH("sp2_step_init_kernel");
  int* hlive = async_steps > 0 ? nullptr : (int*)host_status_slot();
  SEQM_REQUIRE(async_steps > 0 || hlive, "seqm_sp2_density: no page-locked status slot");
  int rc = SEQM_OK;
  auto check = [&](const char* what) {
    cudaError_t e_ = cudaGetLastError();
    ++g_launches;
    if (e_ != cudaSuccess) { set_error(std::string(what) + ": " + cudaGetErrorString(e_)); rc = SEQM_ECUDA; }
  };
  // one purification step of the sharded form: per sub-chunk  compute (stream st) -> all-gather -> unpack
  // (side stream), so the exchange of a sub-chunk overlaps the multiplication of the next one
  auto sharded_step = [&](int step, int* live_counter) {
    cudaStream_t side = cm ? cm->side : st;
    for (int sub = 0; sub < nsub && rc == SEQM_OK; ++sub) {
      double* subrec = records + (size_t)sub * nranks * seg;
      const int r0 = cm ? myrank : 0, r1 = cm ? myrank + 1 : nranks;
      for (int r = r0; r < r1 && rc == SEQM_OK; ++r) {
        Sp2Peers one{};
        one.n = 1;
        one.p[0] = subrec + (size_t)r * seg;
        sp2_step_kernel<true><<<dim3(tps, B), 256, 0, st>>>(B, 0, ld, T, step, eps, dims, occ, X0, X1, stride, state,
                                                          partial, live_counter, r * tpr + sub * tps,
                                                          (sub == 0 && r == r0) ? 1 : 0, one, 0);
        check("sp2_step_kernel<sharded>");
      }
      if (rc) break;
      if (cm) {
        cudaEventRecord(cm->computed[sub], st);
        cudaStreamWaitEvent(side, cm->computed[sub], 0);
        if (nranks > 1) {
          const int r = api->AllGather(subrec + (size_t)myrank * seg, subrec, seg, NCCL_FLOAT64, cm->comm, side);
          if (r) { set_error(std::string("ncclAllGather: ") + api->GetErrorString(r)); rc = SEQM_ECUDA; break; }
        }
      }
      for (int r = 0; r < nranks && rc == SEQM_OK; ++r) {
        sp2_unpack_kernel<<<dim3(tps, B), 256, 0, side>>>(B, ld, T, step, dims, X0, X1, stride, state, partial,
                                                         r * tpr + sub * tps, subrec + (size_t)r * seg, 0);
        check("sp2_unpack_kernel");
      }
    }
    if (cm && rc == SEQM_OK) {
      cudaEventRecord(cm->gathered, side);
      cudaStreamWaitEvent(st, cm->gathered, 0);
    }
  };
  // peer-to-peer form (default with a communicator): ONE launch multiplies all tiles of this rank and stores every
  // new record straight into the record buffer of every rank over NVLink (the exchange overlaps the multiplication
  // tile by tile, no collective moves data); an 8-byte all-gather is the barrier, then every rank unpacks
  const bool p2p = want_p2p && sp2_p2p_ready(cm, (size_t)nranks * seg_full(B, tpr));
  auto p2p_step = [&](int step, int* live_counter) {
    const size_t segf = seg_full(B, tpr);
    Sp2Peers peers{};
    peers.n = nranks;
    for (int q = 0; q < nranks; ++q) peers.p[q] = cm->rec_peer[q] + (size_t)(step & 1) * cm->rec_cap;
    sp2_step_kernel<true><<<dim3(tpr, B), 256, 0, st>>>(B, 0, ld, T, step, eps, dims, occ, X0, X1, stride, state, partial,
                                                      live_counter, myrank * tpr, 1, peers, (size_t)myrank * segf);
    check("sp2_step_kernel<p2p>");
    if (rc) return;
    double* bar = cm->bar + (size_t)(step & 1) * nranks;
    const int r = api->AllGather(bar + myrank, bar, 1, NCCL_FLOAT64, cm->comm, st);
    if (r) { set_error(std::string("ncclAllGather (barrier): ") + api->GetErrorString(r)); rc = SEQM_ECUDA; return; }
    const double* mine = cm->rec_own + (size_t)(step & 1) * cm->rec_cap;
    sp2_unpack_kernel<<<dim3(nranks * tpr, B), 256, 0, st>>>(B, ld, T, step, dims, X0, X1, stride, state, partial, 0, mine, tpr);
    check("sp2_unpack_kernel");
  };
  if (cm) {      // the side stream must not run ahead of the caller's stream
    cudaEventRecord(cm->ready, st);
    cudaStreamWaitEvent(cm->side, cm->ready, 0);
  }
  if (p2p) {
    // entry barrier: a peer may still be unpacking the last step of the PREVIOUS call out of the record buffer
    // this call's first launch is about to store into (cm->bar: [0, 2n) step barriers, [12n, 13n) this one)
    double* eb = cm->bar + (size_t)12 * nranks;
    const int r = api->AllGather(eb + myrank, eb, 1, NCCL_FLOAT64, cm->comm, st);
    if (r) { set_error(std::string("ncclAllGather (entry barrier): ") + api->GetErrorString(r)); return SEQM_ECUDA; }
  }
  // Steps are enqueued in batches; step `it` counts the molecules it leaves live in nlive[it % SP2_NCOUNT] and is a
  // no-op for finished molecules (their CTAs exit on the live flag), so running past convergence is harmless.
  auto enqueue = [&](int it) {
    ++g_sp2_steps;
    int* counter = nlive + (it % SP2_NCOUNT);
    if (p2p) { p2p_step(it, counter); return; }
    if (sharded) { sharded_step(it, counter); return; }
    for (int b0 = 0; b0 < B && rc == SEQM_OK; b0 += 65535) {
      const int nb = B - b0 < 65535 ? B - b0 : 65535;
      sp2_step_kernel<false><<<dim3(ntile, nb), 256, 0, st>>>(B, b0, ld, T, it, eps, dims, occ, X0, X1, stride, state,
                                                             partial, counter, 0, 1, Sp2Peers{}, 0);
      check("sp2_step_kernel");
    }
  };
  if (async_steps > 0) {
    const int nst = (async_steps + 1) & ~1;
    SEQM_REQUIRE(nst <= SP2_NCOUNT, "seqm_sp2_density_async: at most 64 steps");
    prof_mark("sp2_step", st);
    cudaMemsetAsync(nlive, 0, SP2_NCOUNT * sizeof(int), st);
    for (int it = 0; it < nst && rc == SEQM_OK; ++it) enqueue(it);
    if (rc) return rc;
    if (d_unfinished) {
      sp2_accumulate_kernel<<<1, 1, 0, st>>>(nlive + (nst - 1), d_unfinished);
      SEQM_CHECK_LAUNCH("sp2_accumulate_kernel");
    }
  } else {
    int done = 0;
    int batch = g_sp2_hint.load();
    if (batch < 2) batch = 2;
    if (batch > SP2_NCOUNT) batch = SP2_NCOUNT;
    batch = (batch + 1) & ~1;
    for (int it = 0; it < 200 && rc == SEQM_OK && !done;) {
      prof_mark("sp2_step", st);
      cudaMemsetAsync(nlive, 0, SP2_NCOUNT * sizeof(int), st);
      const int it0 = it;
      for (int h = 0; h < batch && rc == SEQM_OK; ++h, ++it) enqueue(it);
      if (rc) break;
      cudaMemcpyAsync(hlive, nlive, SP2_NCOUNT * sizeof(int), cudaMemcpyDeviceToHost, st);
      cudaError_t e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) { set_error(std::string("sp2 loop: ") + cudaGetErrorString(e)); rc = SEQM_ECUDA; break; }
      for (int q = it0; q < it; ++q)
        if (hlive[q % SP2_NCOUNT] == 0) { done = q + 1; break; }
      batch = 2;
    }
    if (done) g_sp2_hint.store((done + 1) & ~1);
  }
  if (rc) return rc;
  int fchunks = (int)(((long long)mmax * mmax + 255) / 256);
  if (fchunks > 512) fchunks = 512;
  prof_mark("sp2_finish", st);
  for (int b0 = 0; b0 < B; b0 += 65535) {
    const int nb = B - b0 < 65535 ? B - b0 : 65535;
    sp2_finish2_kernel<<<dim3(fchunks, nb), 256, 0, st>>>(S, ld, X0, X1, stride, nheavy, dims, active, state, state + B, b0,
                                                          D, niter);
    SEQM_CHECK_LAUNCH("sp2_finish2_kernel");
  }
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_sp2_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, double eps, const uint8_t* active, double* D,
                     int32_t* niter, void* work, void* stream) {
  return sp2_run(nullptr, 0, nmol, molsize, mmax, F, nheavy, nhydro, nocc, eps, active, D, niter, work, stream);
}

unsigned long long seqm_sp2_step_count(void) { return g_sp2_steps.load(); }

int seqm_sp2_density_async(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                           const int32_t* nhydro, const int32_t* nocc, double eps, const uint8_t* active, double* D,
                           int32_t* niter, void* work, int steps, int32_t* unfinished, void* stream) {
  SEQM_REQUIRE(steps > 0, "seqm_sp2_density_async: steps");
  return sp2_run(nullptr, 0, nmol, molsize, mmax, F, nheavy, nhydro, nocc, eps, active, D, niter, work, stream, steps,
                 unfinished);
}

int seqm_sp2_density_nccl(void* comm, int emulate_ranks, int64_t nmol, int64_t molsize, int mmax, const double* F,
                          const int32_t* nheavy, const int32_t* nhydro, const int32_t* nocc, double eps,
                          const uint8_t* active, double* D, int32_t* niter, void* work, void* stream) {
  SeqmComm* cm = (SeqmComm*)comm;
  const int nranks = cm ? cm->nranks : emulate_ranks;
  SEQM_REQUIRE(nranks >= 1 && nranks <= 64, "seqm_sp2_density_nccl: number of ranks");
  return sp2_run(cm, nranks, nmol, molsize, mmax, F, nheavy, nhydro, nocc, eps, active, D, niter, work, stream);
}

}  // extern "C"
