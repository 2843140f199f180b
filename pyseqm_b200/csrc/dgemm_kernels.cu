// Batched FP64 matrix product on the FP64 tensor pipe (DMMA, mma.sync.m8n8k4.f64; tcgen05 has no f64
// kind, so this is the tensor path Blackwell offers for doubles).  C[b] = A[b] * B[b] for square
// row-major matrices; 64x64 CTA tile, 16-deep k slices staged with cp.async double buffering, padded
// shared-memory rows so fragment loads are bank-conflict free.  Used by SP2 purification
// (reference SP2.py:47) and by the Pulay commutator F P - P F (reference scf_loop.py:413-417).
#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int LDA_S = BK + 4;   // 20 doubles: rows land on distinct banks for the A fragment
constexpr int LDB_S = BN + 4;   // 68 doubles

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int bytes = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// grid: (tiles_n, tiles_m, batch); 256 threads = 8 warps as 4 (M) x 2 (N); warp tile 16 x 32
// epilogue: C = alpha*acc (+ beta*Cin if beta != 0), optional per-batch sum of squares / trace
__global__ void __launch_bounds__(256)
dgemm_batched_kernel(int n, int ld, const double* __restrict__ A, const double* __restrict__ B,
                     double* __restrict__ C, long long stride, const uint8_t* __restrict__ active,
                     const int32_t* __restrict__ dims) {
  const int b = blockIdx.z;
  if (active && !active[b]) return;
  const int nn = dims ? dims[b] : n;      // per-molecule size (ragged batches): rows/cols >= nn are skipped
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (m0 >= nn || n0 >= nn) return;
  const double* Ab = A + (size_t)b * stride;
  const double* Bb = B + (size_t)b * stride;
  double* Cb = C + (size_t)b * stride;
  __shared__ __align__(16) double As[2][BM * LDA_S];
  __shared__ __align__(16) double Bs[2][BK * LDB_S];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 1) * 16, wn = (warp & 1) * 32;
  const int g = lane >> 2, t = lane & 3;
  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  auto load_tiles = [&](int buf, int k0) {
    // A tile: 64 rows x 16 cols -> 512 double2, 2 per thread
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256;
      const int r = idx >> 3, c2 = (idx & 7) * 2;
      const int gr = m0 + r, gc = k0 + c2;
      const bool ok = gr < nn && gc < nn;
      cp_async16(&As[buf][r * LDA_S + c2], Ab + (size_t)(ok ? gr : 0) * ld + (ok ? gc : 0), ok);
    }
    // B tile: 16 rows x 64 cols
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256;
      const int r = idx >> 5, c2 = (idx & 31) * 2;
      const int gr = k0 + r, gc = n0 + c2;
      const bool ok = gr < nn && gc < nn;
      cp_async16(&Bs[buf][r * LDB_S + c2], Bb + (size_t)(ok ? gr : 0) * ld + (ok ? gc : 0), ok);
    }
    cp_async_commit();
  };

  const int nk = (nn + BK - 1) / BK;
  load_tiles(0, 0);
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) {
      load_tiles(buf ^ 1, (kt + 1) * BK);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const double* as = As[buf];
    const double* bs = Bs[buf];
#pragma unroll
    for (int kk = 0; kk < BK; kk += 4) {
      double af[2], bf[4];
#pragma unroll
      for (int i = 0; i < 2; ++i) af[i] = as[(wm + i * 8 + g) * LDA_S + kk + t];
#pragma unroll
      for (int j = 0; j < 4; ++j) bf[j] = bs[(kk + t) * LDB_S + wn + j * 8 + g];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = m0 + wm + i * 8 + g;
    if (r >= nn) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + wn + j * 8 + 2 * t;
      if (c + 1 < nn) {
        *reinterpret_cast<double2*>(Cb + (size_t)r * ld + c) = make_double2(acc[i][j][0], acc[i][j][1]);
      } else if (c < nn) {
        Cb[(size_t)r * ld + c] = acc[i][j][0];
      }
    }
  }
}

int launch_dgemm_batched(int n, int ld, const double* A, const double* B, double* C, long long stride, int batch,
                         const uint8_t* active, const int32_t* dims, cudaStream_t st) {
  if (batch <= 0 || n <= 0) return SEQM_OK;
  if (ld % 2 != 0) return fail(SEQM_EINVAL, "dgemm: leading dimension must be even (16-byte rows)");
  if (batch > 65535) {
    // gridDim.z limit: split the batch
    for (int b0 = 0; b0 < batch; b0 += 65535) {
      const int nb = batch - b0 < 65535 ? batch - b0 : 65535;
      int rc = launch_dgemm_batched(n, ld, A + (size_t)b0 * stride, B + (size_t)b0 * stride, C + (size_t)b0 * stride,
                                    stride, nb, active ? active + b0 : nullptr, dims ? dims + b0 : nullptr, st);
      if (rc) return rc;
    }
    return SEQM_OK;
  }
  dim3 grid(cdiv(n, BN), cdiv(n, BM), batch);
  prof_mark("dgemm", st);
  dgemm_batched_kernel<<<grid, 256, 0, st>>>(n, ld, A, B, C, stride, active, dims);
  SEQM_CHECK_LAUNCH("dgemm_batched_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

}  // namespace seqm

using namespace seqm;

extern "C" int seqm_dgemm_batched(int64_t n, int64_t ld, const double* A, const double* B, double* C,
                                  int64_t stride, int64_t batch, const uint8_t* active, void* stream) {
  SEQM_REQUIRE(A && B && C && n >= 0 && ld >= n && stride >= 0, "seqm_dgemm_batched: bad argument");
  return launch_dgemm_batched((int)n, (int)ld, A, B, C, stride, (int)batch, active, nullptr, (cudaStream_t)stream);
}
