/*
 * seqm_b200 — C ABI of the B200 (sm_100a) kernels behind PYSEQM's batched SCF hot path.
 *
 * The reference (mtzgroup/PYSEQM) is pure Python/PyTorch and has no FFI of its own; the boundary it
 * offers is the Python call surface of seqm/basics.py and seqm/seqm_functions/*.py.  Each entry point
 * below replaces the body of one reference function (cited), is called from pyseqm_b200/_ops.py through
 * ctypes, and is what a maintainer of the reference would bind instead of the eager-ATen code
 * (INTEGRATION.md shows the stub).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to caller-owned contiguous memory (PyTorch allocations); scratch is
 *     passed in (seqm_*_workspace).  The library itself allocates only (i) 512 bytes of page-locked host memory
 *     per calling thread (the status word the host-driven loops read back) and (ii), for the multi-GPU SP2 form,
 *     the peer-mapped record buffers of a communicator (seqm_nccl_comm_create .. seqm_nccl_comm_destroy).
 *   - all floating point data is fp64, index data is int32 unless noted (the int64 index tensors of
 *     the reference API are narrowed once per geometry by the host layer).
 *   - calls are asynchronous on `stream` (a cudaStream_t passed as void*) and do not synchronise unless
 *     documented.  Entry points may be called concurrently from several host threads on different streams
 *     (the launch counter is atomic, the profiler's mark list is locked, error text is thread local).
 *     Process-wide state is limited to the TEST / TUNING hooks marked "process-wide" below (Jacobi tolerance and
 *     variant selection, fused-SCF switch, adjoint mode, SP2 sharding communicator): set them before the threads
 *     start; everything a caller needs per call travels in the arguments (seqm_scf_options_t).
 *   - return value: 0 on success, negative seqm_status_t otherwise; seqm_last_error() gives the text
 *     (thread local).
 *
 * Layouts
 *   padded matrices   X[B][n][n], n = 4*S, atom slot a owns rows/cols 4a..4a+3 = (s,px,py,pz)
 *   block matrices    M[B*S*S][4][4], block index b*S*S + a*S + c  (reference hcore.py:73-158)
 *   two-centre ints   stored per pair CLASS (HH, XH, XX: hydrogen/heavy), tiles of 32 pairs with the
 *                     pair index fastest: wXX[tile][100][32], wXH[tile][10][32], wHH[pair]; the packed
 *                     10x10 index is the reference's (k on atom i)*10 + (l on atom j), fock.py:76-91.
 *                     seqm_w_expand() materialises the reference's dense w[P][10][10].
 */
#ifndef SEQM_B200_H
#define SEQM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  SEQM_OK = 0,
  SEQM_EINVAL = -1,      /* bad argument (null pointer, negative size, unsupported size) */
  SEQM_EUNSUPPORTED = -2,/* element / method / size outside the supported range */
  SEQM_ECUDA = -3        /* a CUDA runtime call or launch failed */
} seqm_status_t;

#define SEQM_CLS_HH 0
#define SEQM_CLS_XH 1
#define SEQM_CLS_XX 2

/* Pair/atom topology of one batch: the Parser's index tensors (reference basics.py:43-118) narrowed to
 * int32 plus the CSR tables the no-atomics reductions use. */
typedef struct {
  int64_t nmol, molsize, natoms, npairs;
  const int32_t* Z;          /* [A] atomic number */
  const int32_t* atom_mol;   /* [A] molecule of the atom */
  const int32_t* atom_slot;  /* [A] slot in its molecule: maskd = mol*S*S + slot*(S+1) */
  const int32_t* idxi;       /* [P] first atom of the pair (index into the real-atom list) */
  const int32_t* idxj;       /* [P] second atom */
  const int32_t* pair_cls;   /* [P] SEQM_CLS_* */
  const int32_t* pair_slot;  /* [P] position of the pair inside its class list */
  const int32_t* cls_pairs[3]; /* [ncls[c]] pair ids of each class, ascending */
  int64_t ncls[3];
  const int32_t* start_i;    /* [A+1] pairs whose first atom is a: [start_i[a], start_i[a+1]) */
  const int32_t* start_j;    /* [A+1] range in perm_j of the pairs whose second atom is a */
  const int32_t* perm_j;     /* [P] pair ids ordered by second atom */
} seqm_topology_t;

/* Two-centre integrals in class-tiled storage. */
typedef struct {
  double* wHH; /* [ncls[HH]] */
  double* wXH; /* [ceil(ncls[XH]/32)][10][32] */
  double* wXX; /* [ceil(ncls[XX]/32)][100][32] */
} seqm_w_t;

const char* seqm_last_error(void);
/* kernels launched by this library since load (bench.py's gpu_launches) */
unsigned long long seqm_launch_count(void);
/* optional CUDA-event profiler: regions between consecutive launches groups; report() synchronises and writes
 * "name total_ms count" lines */
void seqm_profile_enable(int on);
int seqm_profile_report(char* buf, size_t cap);
int seqm_version(void);
size_t seqm_atompar_bytes(void); /* size of one packed per-atom record (seqm_atom_pack output) */

/* Per-atom parameter record: Slater exponents, beta, the point-charge multipole parameters
 * dd, qq, rho0, rho1, rho2 (reference two_elec_two_center_int.py:7-52, cal_par.py:8-173), core-repulsion
 * parameters.  gK/gL/gM are [A][ngauss] row-major or NULL when ngauss == 0. */
int seqm_atom_pack(int64_t natoms, const int32_t* Z, const double* zeta_s, const double* zeta_p,
                   const double* beta_s, const double* beta_p, const double* g_ss, const double* g_pp,
                   const double* g_p2, const double* h_sp, const double* alpha, const double* gK,
                   const double* gL, const double* gM, int ngauss, void* atompar_out, void* stream);

/* Pair geometry of the Parser (reference basics.py:91-108) on the device: R[B][S][3] (Angstrom) ->
 * xij[P][3] = (R_j - R_i)/d, rij[P] = d*lcf (zero distance^2 -> 1e-4), and its adjoint:
 * gR[B][S][3] (must be zero filled: padding atoms are not written) from gx[P][3], gr[P]. */
int seqm_pair_geometry(const seqm_topology_t* topo, const double* R, double lcf, double* xij, double* rij,
                       void* stream);
int seqm_pair_geometry_grad(const seqm_topology_t* topo, const double* xij, const double* rij, const double* gx,
                            const double* gr, double lcf, double* gR, void* stream);

/* Parser.forward (reference basics.py:43-118) on the device, with the outer cutoff: the pair list in the reference's
 * order (molecule, a, b > a), its int64 index tensors and geometry, and the tables of seqm_topology_t, from one pass
 * over the atom slots.  species[B][S] int64, R[B][S][3] in Angstrom.
 *   seqm_parser_count  counts5[5][B*S] (int32): first-atom pairs of class HH / XH / XX, second-atom pairs, real-atom flag
 *   (the caller scans the five rows exclusively into int64 offsets: off_pair = scan of the class total)
 *   seqm_parser_fill   writes everything at those offsets; pairid_scratch: B*S*S int32 */
int seqm_parser_count(int64_t nmol, int64_t molsize, const int64_t* species, const double* R, double cutoff,
                      double lcf, int32_t* counts5, void* stream);
int seqm_parser_fill(int64_t nmol, int64_t molsize, const int64_t* species, const double* R, double cutoff, double lcf,
                     const int64_t* off_pair, const int64_t* off_hh, const int64_t* off_xh, const int64_t* off_xx,
                     const int64_t* off_second, const int64_t* off_real, int64_t* mask, int64_t* pair_molid,
                     int64_t* ni, int64_t* nj, int64_t* idxi64, int64_t* idxj64, double* xij, double* rij,
                     int32_t* idxi, int32_t* idxj, int32_t* pair_cls, int32_t* pair_slot, int32_t* cls_hh,
                     int32_t* cls_xh, int32_t* cls_xx, int32_t* perm_j, int32_t* pairid_scratch, void* stream);

/* hcore() (reference hcore.py:6-167) = two-centre integrals + rotate + diatomic overlap + core
 * Hamiltonian, and pair_nuclear_energy() (energy.py:38-78).
 *   xij[P][3], rij[P]      Parser geometry (unit vector i->j, distance in bohr)
 *   w                      class-tiled integrals (out)
 *   M[B*S*S][16]           block core Hamiltonian (out; must be zero-filled by the caller: only the upper
 *                          triangle blocks / upper triangle of diagonal blocks are written)
 *   EnucAB[P]              core-core repulsion per pair (out, may be NULL)
 *   pairvec[P][20]         scratch */
int seqm_hcore(const seqm_topology_t* topo, const void* atompar, const double* xij, const double* rij,
               const double* U_ss, const double* U_pp, int ngauss, seqm_w_t w, double* M, double* EnucAB,
               double* pairvec, void* stream);

/* dense w[P][10][10] of the reference API from the class-tiled storage */
int seqm_w_expand(const seqm_topology_t* topo, seqm_w_t w, double* w_dense, void* stream);

/* fock() (reference fock.py:6-139) and elec_energy() (energy.py:19-36).
 *   P[B][n][n] density, M block core Hamiltonian, F[B][n][n] (out, full symmetric; must have been zero
 *   filled once by the caller: entries outside the pair pattern are never written)
 *   active[B] (may be NULL): molecules with active[b]==0 are skipped (frozen once converged)
 *   Eelec[B] (may be NULL): 0.5*sum P.(Hcore+F) for the active molecules
 *   work (may be NULL): seqm_elec_energy_workspace() bytes; with it the energy sum of a molecule is cut into
 *   chunks of 8192 matrix elements (one CTA each, partial sums added in a fixed order) instead of one CTA per
 *   molecule - needed when the batch is a few large molecules (config 5 is one) */
int seqm_fock(const seqm_topology_t* topo, seqm_w_t w, const double* P, const double* M,
              const double* g_ss, const double* g_pp, const double* g_sp, const double* g_p2,
              const double* h_sp, const uint8_t* active, double* F, double* Eelec, double* pairvec,
              void* work, void* stream);

/* elec_energy() alone: E[b] = 0.5 * sum_ij P_ij (H_ij + F_ij) with H the symmetrised block matrix M.
 * xl != 0: XL-BOMD shadow energy  sum D.F - 0.5 (F-H).P  (reference XLBOMD.py:40-52), D given. */
size_t seqm_elec_energy_workspace(int64_t nmol, int64_t molsize);
int seqm_elec_energy(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                     const double* D, int xl, const uint8_t* active, double* E, void* work, void* stream);

/* seqm_elec_energy for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and arithmetic as the chunked form of seqm_elec_energy (per-chunk partial sums of 8192 matrix
 * elements, added in a fixed order), but the partial sums live in a file-scope `static __device__` array of
 * SEQM_NOALLOC_MAX_PARTIALS doubles (512 KiB, held by every GPU that loads the library until exit) instead of caller
 * scratch: the call allocates nothing - no cudaMalloc / cudaFree, no allocating library routine - and takes no
 * workspace argument.  Limit: nmol * ceil(16 * molsize^2 / 8192) <= SEQM_NOALLOC_MAX_PARTIALS, otherwise SEQM_EINVAL with
 * a message that names the sizes (e.g. 65,536 molecules of up to 22 atoms, or one system of up to 5,792 atoms).
 * The array is shared by every call on a device and keeps the last call's contents; nothing in it needs clearing (a call
 * only reads partial sums it has written itself), but two calls that OVERLAP on the GPU (different streams) race on it:
 * serialise them.  The first call on a device resolves the array's device address (cudaGetSymbolAddress, once per
 * device with that device current): make that first call outside stream capture. */
#define SEQM_NOALLOC_MAX_PARTIALS 65536
int seqm_elec_energy_noalloc(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                             const double* D, int xl, const uint8_t* active, double* E, void* stream);

/* sym_eig_trunc() (reference diag.py:102-172) with pack/unpack (pack.py:7-41) folded in: batched
 * symmetric eigensolver on the packed matrix (heavy atoms 4 orbitals, hydrogens 1) and
 * P = 2 C_occ C_occ^T written back in padded layout.  One CTA per molecule, cyclic Jacobi in shared
 * memory; supports packed size m <= seqm_eigh_max_m().
 *   nheavy,nhydro,nocc [B] int32;  e_out[B][n] (may be NULL) ascending orbital energies, zero padded;
 *   v_out[B][n][n] (may be NULL) eigenvectors in the reference's packed layout C[packed orbital][k];
 *   info[B] (may be NULL) sweeps used.  mmax: largest packed size in the batch (sizes shared memory). */
int seqm_eigh_max_m(void);
void seqm_eigh_force_block(int on); /* process-wide test hook: shared-memory (CTA) Jacobi also for m <= 32 */
void seqm_eigh_warp_variant(int ring); /* process-wide test hook: warp kernels (m <= 32, stand-alone and inside the one-launch SCF): 1 (default) = one-sided ring Jacobi in registers (csrc/eigh_ring.cuh), 0 = two-sided odd-even Jacobi (csrc/eigh_warp.cuh) */
void seqm_eigh_block_config(int bits); /* process-wide tuning hook: shared-memory Jacobi variants for 32 < m <= 64: bit 0 = 256-thread CTAs (4 per SM) for m <= 48, bit 1 = FP32 rotation tangent (FP64 cosine) for m <= 64; default 3 (measured 135 -> 118 ms on the m > 32 molecules of config 4, tools/exp_organics_big.py) */
void seqm_eigh_tolerance(double tol); /* process-wide: Jacobi stopping rule, one final sweep once off^2 <= tol * total^2 (default 1e-16 for both kernels - r02a: 1e-14 fails the 1e-9 density check on one fixture, 1e-16 passes all and is 4.6 % faster than 1e-18; tol <= 0 restores it) */
/* process-wide, like the reference's module-level diag.CHECK_DEGENERACY (diag.py:7-10,79-98): orbitals within 1e-14 of
 * the HOMO share the remaining electrons equally.  Implemented in the warp kernels (packed size <= 32, stand-alone and
 * inside the one-launch SCF); the host layer raises for larger molecules. */
void seqm_eigh_check_degeneracy(int on);
int seqm_eigh_check_degeneracy_max_m(void);
int seqm_eigh_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                      const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P,
                      double* e_out, double* v_out, int32_t* info, void* stream);

/* SP2() (reference SP2.py:3-72) on packed matrices with pack/unpack folded in.  Fpad[B][n][n] in,
 * D[B][n][n] out (= 2 X, padded layout).  work: seqm_sp2_workspace(...) bytes of scratch.
 * niter[B] (may be NULL) multiplications used.  Synchronises the stream: the host reads the per-step counters of
 * live molecules, the first time after as many steps as the previous call needed (one synchronisation per call in a
 * steady-state loop such as MD or the later SCF iterations), then every two steps.
 * seqm_sp2_density_async: exactly `steps` (<= 64, rounded up to even) purification steps are enqueued, finished
 * molecules are skipped by their CTAs; no host synchronisation or allocation, so the call can be captured in a CUDA
 * graph; the number of molecules still live after the last step is ADDED to unfinished[0] (DEVICE int32, a running
 * total the caller zeroes and inspects when convenient). */
size_t seqm_sp2_workspace(int64_t nmol, int64_t molsize);
int seqm_sp2_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, double eps, const uint8_t* active,
                     double* D, int32_t* niter, void* work, void* stream);
unsigned long long seqm_sp2_step_count(void); /* purification steps enqueued since load (bench.py: time per step) */
int seqm_sp2_density_async(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                           const int32_t* nhydro, const int32_t* nocc, double eps, const uint8_t* active, double* D,
                           int32_t* niter, void* work, int steps, int32_t* unfinished, void* stream);

/* Single large system (BASELINE.json configs[4]): the same purification with the 64x64 tiles of the upper block
 * triangle of X^2 SHARDED over the GPUs of one NCCL communicator (one process per GPU).  Rank g multiplies the
 * tiles [g*tpr, (g+1)*tpr) of the enumeration, writes each new tile and its two partial sums (diagonal, squared
 * Frobenius norm) into a contiguous record, the records are all-gathered over NVLink (ncclAllGather, in place) and
 * every rank unpacks all tiles and their transposes into its own copy of X; the trace rule is evaluated by every
 * rank from the same gathered partial sums in the same order, so all ranks take identical branches and the result
 * is bit-identical to seqm_sp2_density.  The range of a rank is cut into pieces (seqm_sp2_nccl_subchunks(n), n = 0:
 * automatic - as many, up to 4, as still leave 148 x 4 tiles per launch): the all-gather + unpack of piece k run on a
 * side stream while piece k+1 is multiplied.
 *   NCCL is bound at run time (dlopen of the libnccl.so.2 already in the process); the library owns the
 *   communicator it creates:  rank 0: seqm_nccl_unique_id(id) -> broadcast the 128 bytes by any means ->
 *   every rank: seqm_nccl_comm_create(id, rank, nranks, &comm) with its CUDA device current.
 *   comm == NULL with emulate_ranks >= 1 runs the tile ranges of that many ranks one after the other on this GPU
 *   through the same records and unpack kernel (test form, no NCCL).
 *   seqm_scf_sp2_comm(comm, emulate_ranks) makes the SP2 inside seqm_scf use the sharded form (NULL, 0 restores the
 *   single-GPU form); replaces the role of torch.matmul in reference SP2.py:47 for one large system.
 *   Default with a communicator of <= 16 ranks: the PEER-TO-PEER form - the library allocates the record buffers
 *   itself, maps every rank's buffer into every process (CUDA IPC handles exchanged through the communicator) and
 *   the multiplication kernel stores each finished record directly into all ranks' buffers over NVLink, so the
 *   exchange overlaps the multiplication tile by tile inside ONE launch; an 8-byte all-gather serves as the barrier
 *   before the unpack.  seqm_sp2_nccl_p2p(0), more than 16 ranks or a failed IPC mapping select the ncclAllGather
 *   form described above. */
void seqm_sp2_nccl_p2p(int on);
int seqm_nccl_unique_id(void* id128);
int seqm_nccl_comm_create(const void* id128, int rank, int nranks, void** comm_out);
int seqm_nccl_comm_destroy(void* comm);
void seqm_sp2_nccl_subchunks(int n);
void seqm_sp2_shard_plan(int ntile, int nranks, int nsub, int rank, int32_t* out4); /* host only: {tiles per rank, tiles per piece, first owned tile, end} */
size_t seqm_sp2_nccl_workspace(int64_t nmol, int64_t molsize, int nranks);
int seqm_sp2_density_nccl(void* comm, int emulate_ranks, int64_t nmol, int64_t molsize, int mmax, const double* F,
                          const int32_t* nheavy, const int32_t* nhydro, const int32_t* nocc, double eps,
                          const uint8_t* active, double* D, int32_t* niter, void* work, void* stream);
void seqm_scf_sp2_comm(void* comm, int emulate_ranks);

/* The SCF driver: scf_forward0/1/2 (reference scf_loop.py:32-510) run device resident.
 *   converger: 0 constant mixing (alpha), 1 adaptive, 2 adaptive then Pulay
 *   P[B][n][n] in: initial guess, out: converged density; F out: fock(P) at the final density
 *   notconverged[B] out; niter[B] out (iterations taken by each molecule)
 *   iterations_out: int32[2] on the HOST = {iterations run, molecules still unconverged}
 *   work: seqm_scf_workspace(...) bytes.
 * Two execution forms, same arithmetic: batches whose largest packed size is <= 32 with the eigh density and
 * debug == 0 run the whole loop in ONE kernel (one warp per molecule, one stream synchronisation at the end);
 * otherwise the host drives one iteration at a time and synchronises once per iteration (convergence count),
 * like the reference's notconverged.any().  seqm_scf_fused(0) forces the second form (test hook). */
typedef struct {
  int converger;      /* 0, 1, 2 */
  double alpha;       /* converger 0 */
  int use_sp2;        /* density from SP2 instead of eigh */
  double sp2_eps;
  double eps;         /* |dEelec| threshold, eV */
  int max_iter;
  int debug;          /* print "scf k max_err n_notconverged" like scf_loop.debug */
  int async_status;   /* one-launch form only: do not synchronise; iterations_out then points to 16 bytes of
                         page-locked host memory that receive two uint64 {molecules left, iterations run} when the
                         stream reaches the end of the SCF */
} seqm_scf_options_t;
void seqm_scf_fused(int on);
size_t seqm_scf_workspace(int64_t nmol, int64_t molsize, const seqm_scf_options_t* opt);
int seqm_scf(const seqm_topology_t* topo, seqm_w_t w, const double* M, const double* g_ss,
             const double* g_pp, const double* g_sp, const double* g_p2, const double* h_sp,
             const int32_t* nheavy, const int32_t* nhydro, const int32_t* nocc, int mmax,
             const seqm_scf_options_t* opt, double* P, double* F, double* Eelec, uint8_t* notconverged,
             int32_t* niter, double* pairvec, void* work, int32_t* iterations_out, void* stream);

/* Adjoint of hcore + fock + elec_energy + pair_nuclear_energy at fixed density (what autograd replays in
 * the reference: basics.py:369-388, XLBOMD.py:203-217).
 *   E = sum_b gE[b]*Eelec_b + sum_p gNuc[p]*EnucAB_p ;  D == NULL: Eelec = 0.5 sum P.(H+F(P)),
 *   D != NULL (XL-BOMD): Eelec = sum D.F(P) - 0.5 (F(P)-H).P.
 *   out: grad_xij[P][3], grad_rij[P]  (dE/dxij, dE/drij; chained through the Parser by autograd) */
void seqm_pair_grad_reverse(int on); /* heavy-heavy pairs: 1 (default) reverse-mode kernel, 0 forward-mode duals (comparison / test hook) */
int seqm_pair_grad(const seqm_topology_t* topo, const void* atompar, const double* xij, const double* rij,
                   int ngauss, const double* P, const double* D, const double* gE, const double* gNuc,
                   double* grad_xij, double* grad_rij, void* stream);

/* C[b] = A[b] * B[b], square row-major fp64 matrices on the FP64 tensor pipe (building block of SP2 and of
 * the Pulay commutator; exported for tests).  ld even, stride in elements, active[batch] may be NULL. */
int seqm_dgemm_batched(int64_t n, int64_t ld, const double* A, const double* B, double* C, int64_t stride,
                       int64_t batch, const uint8_t* active, void* stream);

/* The elementwise half of an XL-BOMD / velocity-Verlet step (reference XLBOMD.py:277-307, MolecularDynamics.py:229-273),
 * so that a whole MD step consists of launches of this library and can be captured in one CUDA graph.  Arithmetic in the
 * order of the reference's tensor expressions, without FMA contraction.
 *   seqm_xl_extrapolate   P = kappa D + sum_j coeff2m[cidx + j] Pt[j], Pt[m-1-cidx] = P with cidx = step % m
 *                         (XLBOMD.py:293-295; coeff2m = the 2m doubled coefficients of :259-262, a HOST array);
 *                         step from *step_dev (DEVICE int64, so one captured graph serves every step) or, if NULL,
 *                         step_host.  nelem = elements of one density matrix batch (B n n); Pt[m][nelem].
 *   seqm_verlet_kick_drift  v += 0.5 a dt ; x += v dt                    (n = 3 x atoms incl. padding)
 *   seqm_verlet_finish    force = -gR ; a = force / mass * acc_scale ; v += 0.5 a dt ; (*step_dev)++ if not NULL;
 *                         mass[n/3] per atom slot (1 for padding atoms)
 *   seqm_energy_finish    Enuc[b] = sum of EnucAB over the pairs [pair_off[b], pair_off[b+1]) in a fixed order,
 *                         Etot = Eelec + Enuc, Hf = Etot + hconst[b] (hconst = sum eheat - sum Eiso of the molecule:
 *                         energy.py:92-94,109-118)
 *   seqm_md_thermo        Ek[b], T[b] (MolecularDynamics.py:229-233) and, if P != NULL, atomic charges q[B][S] =
 *                         tore - sum_mu P_mumu and dipole[B][3] = sum q x (:235-243) */
int seqm_xl_extrapolate(int64_t nelem, int m, double kappa, const double* coeff2m, const int64_t* step_dev,
                        int64_t step_host, const double* D, double* Pt, double* P, void* stream);
int seqm_verlet_kick_drift(int64_t n, double dt, const double* acc, double* v, double* x, void* stream);
int seqm_verlet_finish(int64_t n, double dt, double acc_scale, const double* gR, const double* mass, double* force,
                       double* acc, double* v, int64_t* step_dev, void* stream);
int seqm_energy_finish(int64_t nmol, const int32_t* pair_off, const double* EnucAB, const double* Eelec,
                       const double* hconst, double* Enuc, double* Etot, double* Hf, void* stream);
int seqm_md_thermo(int64_t nmol, int64_t molsize, double ke_scale, double ev_to_kelvin, const double* mass,
                   const double* tore, const int32_t* natoms, const double* v, const double* x, const double* P,
                   double* Ek, double* T, double* q, double* dipole, void* stream);

#ifdef __cplusplus
}
#endif
#endif
