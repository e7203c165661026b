/* cyten_b200 -- C-ABI of the B200-native abelian block-sparse hot path.
 *
 * Every entry point is `extern "C"`, takes plain pointers / sizes, returns 0 on success and a
 * non-zero CYB_ERR_* code otherwise (the message is available from cyb_last_error()).
 * Unless stated otherwise every pointer is a DEVICE pointer, every integer table is row-major
 * int64 (the interchange format of the reference's BlockInds,
 * include/cyten/backends/block_inds.h:22-160), and `stream` is a cudaStream_t passed as void*.
 *
 * Block data lives in a per-tensor "arena": one contiguous device allocation; block n is a
 * C-contiguous array starting at element offset `offs[n]` (elements of the arena dtype),
 * which mirrors the reference's AbelianBackendData::blocks (include/cyten/backends/abelian.h:51-83).
 *
 * The reference function each entry point replaces is cited per declaration
 * (paths relative to the tenpy/cyten source tree).
 */
#ifndef CYTEN_B200_H
#define CYTEN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CYB_VERSION 100
#define CYB_MAX_RANK 8

enum cyb_dtype { CYB_F64 = 0, CYB_C128 = 1 };

enum cyb_error {
    CYB_OK = 0,
    CYB_ERR_INVALID = 1,   /* bad argument */
    CYB_ERR_CUDA = 2,      /* a CUDA runtime call failed */
    CYB_ERR_LIMIT = 3,     /* table larger than a built-in limit */
    CYB_ERR_NOCONV = 4     /* iterative factorisation did not converge */
};

typedef void* cyb_stream_t;

/* ---- library ------------------------------------------------------------------------- */
int cyb_version(void);
const char* cyb_last_error(void);
/* sm_count / compute capability of the current device */
int cyb_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- (a1) BlockInds bookkeeping: src/backends/block_inds.cpp --------------------------- */

/* Stable lexsort of the rows of `rows[nrows][ncols]`, LAST column primary
 * (== np.lexsort(rows.T)); replaces BlockInds::lexsort_indices (block_inds.cpp:290-299).
 * perm_out[nrows]: perm_out[r] = index of the row that goes to sorted position r. */
int cyb_lexsort_rows(const int64_t* rows, int64_t nrows, int64_t ncols, int64_t* perm_out,
                     cyb_stream_t stream);

/* out[r][:] = rows[idx[r]][cols[:]]; replaces BlockInds::take / take_columns
 * (block_inds.cpp). `cols` may be NULL (all columns, ncols_out == ncols). */
int cyb_take_rows(const int64_t* rows, int64_t ncols, const int64_t* idx, int64_t nrows_out,
                  const int64_t* cols, int64_t ncols_out, int64_t* out, cyb_stream_t stream);

/* Start index of every run of equal consecutive rows, followed by nrows (include_len=true);
 * replaces BlockInds::find_row_differences (block_inds.cpp:308-336).
 * starts_out[nrows+1], count_out[1] = number of entries written (runs + 1). */
int cyb_row_differences(const int64_t* rows, int64_t nrows, int64_t ncols, int64_t* starts_out,
                        int64_t* count_out, cyb_stream_t stream);

/* keys_out[r] = sum_c rows[r][c] * strides[c]; replaces BlockInds::pack (block_inds.cpp:556-571). */
int cyb_pack_rows(const int64_t* rows, int64_t nrows, int64_t ncols, const int64_t* strides,
                  int64_t* keys_out, cyb_stream_t stream);

/* Merge-join of two ascending, duplicate-free int64 key lists; replaces iter_common_sorted_1d
 * (cyten/tools/misc.py iter_common_sorted, used at abelian.cpp:1430-1433).
 * pairs_out[min(na,nb)][2] receives (i, j) with a[i] == b[j] in ascending order, count_out[1]. */
int cyb_common_sorted(const int64_t* a, int64_t na, const int64_t* b, int64_t nb,
                      int64_t* pairs_out, int64_t* count_out, cyb_stream_t stream);

/* ---- (a3,a4,a5) batched strided block copies ------------------------------------------- *
 * One job copies a rank-`rank` box between two strided views; permute_legs (abelian.cpp:988-1006
 * + BlockBackend::permute_axes), combine_legs (abelian.cpp:1196-1217 zeros + set_item of
 * BlockBackend::combine_legs, block_backend.cpp:784-837) and split_legs (abelian.cpp:3418-3433
 * get_item + reshape + permute_axes) are all expressed as lists of such jobs. Offsets and
 * strides are in ELEMENTS of `dtype`. */
typedef struct cyb_copy_job {
    int64_t src_off;
    int64_t dst_off;
    int64_t rank;
    int64_t shape[CYB_MAX_RANK];
    int64_t src_stride[CYB_MAX_RANK];
    int64_t dst_stride[CYB_MAX_RANK];
} cyb_copy_job;

/* Executes all jobs src -> dst in one launch.  `jobs` is a DEVICE array, `jobs_host` its HOST copy (the
 * library sizes its schedule from it).  `conj` != 0 conjugates complex data. */
int cyb_copy_blocks(const void* src, void* dst, const cyb_copy_job* jobs, const cyb_copy_job* jobs_host,
                    int64_t njobs, int dtype, int conj, cyb_stream_t stream);

/* The two halves of cyb_copy_blocks, for callers that keep a job list (a "plan": the bookkeeping of
 * permute_legs / combine_legs / split_legs depends only on block_inds and the legs, not on the block
 * contents) and replay it on new data.  prepare canonicalises the jobs and resolves them into work
 * units inside `workspace` (device memory, >= cyb_copy_blocks_workspace(jobs_host, njobs) bytes,
 * caller-owned, must stay alive); run is the single copy launch. */
int64_t cyb_copy_blocks_workspace(const cyb_copy_job* jobs_host, int64_t njobs);
int cyb_copy_blocks_prepare(const cyb_copy_job* jobs, int64_t njobs, void* workspace,
                            int64_t workspace_bytes, cyb_stream_t stream);
int cyb_copy_blocks_run(const void* src, void* dst, const void* workspace, int64_t njobs, int dtype,
                        int conj, cyb_stream_t stream);

/* ---- (a6) compose: abelian.cpp:1238-1469 ------------------------------------------------ *
 * Device planner for abelian_compose_worker.  Inputs describe A (block_inds a_bi[na][ra],
 * the first `a_keep` columns are kept, the rest contracted) and B (b_bi[nb][rb], the first
 * `ncontr` columns contracted, the rest kept); both lexsorted as AbelianBackendData keeps them.
 *  contr_nsec[ncontr]        number of sectors of each contracted leg, in B's column order
 *  a_charge[na][nsym]        fused charge of the kept legs of every A row   (host computes the
 *  b_charge[nb][nsym]        per-leg sector lookups into these flat tables on device, see
 *                            cyb_fuse_charges)
 *  a_mk[na][2], b_kn[nb][2]  (M, K) of each A block as a matrix and (K, N) of each B block
 *  a_off[na], b_off[nb]      arena offsets
 * Outputs (device):
 *  header[8]: [0]=number of output blocks, [1]=number of (A,B) block pairs, [2]=total output
 *             elements (aligned), [3]=1 if the result fit into cap_out / cap_pairs (else only the counts are valid),
 *             [4]=sum M*K*N over pairs (algorithmic flops / 2), [5],[6]=number of A row / B column groups
 *  out_bi[<=cap_out][a_keep + rb - ncontr]  lexsorted result block_inds
 *  out_rowcol[<=cap_out][2]   (row_a group representative row in A order, col_b representative)
 *  out_mn[<=cap_out][2], out_off[<=cap_out], out_pair_ptr[<=cap_out + 1]
 *  pairs[<=cap_pairs][3]      (a_off, b_off, K) per matched block pair, grouped by output block
 */
int64_t cyb_compose_plan_workspace(int64_t na, int64_t nb);
int cyb_compose_plan(const int64_t* a_bi, int64_t na, int64_t ra, int64_t a_keep,
                     const int64_t* b_bi, int64_t nb, int64_t rb, int64_t ncontr,
                     const int64_t* contr_nsec, const int64_t* a_charge, const int64_t* b_charge,
                     int64_t nsym, const int64_t* a_mk, const int64_t* b_kn, const int64_t* a_off,
                     const int64_t* b_off, int64_t align_elems, int64_t cap_out, int64_t cap_pairs,
                     int64_t* header, int64_t* out_bi, int64_t* out_rowcol, int64_t* out_mn,
                     int64_t* out_off, int64_t* out_pair_ptr, int64_t* pairs, void* workspace,
                     int64_t workspace_bytes, cyb_stream_t stream);

/* charge[r][s] = fuse_f sectors_f[bi[r][col_f]][s] for the abelian fusion rule (sum, mod
 * modulus[s] when modulus[s] > 0); replaces Symmetry::multiple_fusion_broadcast
 * (src/symmetries/symmetry.cpp:375-402) for U(1)/Z_N products.
 *  sectors: all legs' sector_decomposition tables concatenated, leg f starts at row
 *  sector_ptr[f]; cols[f] is the block_inds column of leg f. */
int cyb_fuse_charges(const int64_t* bi, int64_t nrows, int64_t ncols, const int64_t* cols,
                     int64_t nlegs, const int64_t* sectors, const int64_t* sector_ptr,
                     const int64_t* modulus, int64_t nsym, int64_t* charge_out,
                     cyb_stream_t stream);

/* Grouped GEMM  C_t = sum_{p in pairs of t} A_p (M_t x K_p, row-major, lda = K_p)
 *                                         * B_p (K_p x N_t, row-major, ldb = N_t)
 * for every output block t; replaces the matrix_dot / operator+ loop of
 * abelian.cpp:1424-1450 (NumpyBlockBackend::matrix_dot = np.dot, numpy.cpp:1219-1225).
 * Tables are the outputs of cyb_compose_plan. float64 and complex128 run on DMMA
 * (mma.sync m8n8k4 f64). `nblocks` is the number of output blocks to compute: all of them when
 * `block_subset` is NULL, else the block ids listed in block_subset[nblocks] (a rank's share of
 * the sectors under multi-GPU sharding). */
int cyb_grouped_gemm(const void* a, const void* b, void* c, int dtype, const int64_t* out_mn,
                     const int64_t* out_off, const int64_t* out_pair_ptr, const int64_t* pairs,
                     const int64_t* block_subset, int64_t nblocks, cyb_stream_t stream);

/* The two halves of cyb_grouped_gemm for a cached compose plan: prepare builds the tile schedule
 * (cost-sorted block order, tile prefix sums for the DMMA and the streaming kernel) into `workspace`
 * (>= cyb_grouped_gemm_workspace(nblocks) bytes, caller-owned); run launches the kernels. */
int64_t cyb_grouped_gemm_workspace(int64_t nblocks);
int cyb_grouped_gemm_prepare(int dtype, const int64_t* out_mn, const int64_t* out_pair_ptr,
                             const int64_t* pairs, const int64_t* block_subset, int64_t nblocks,
                             void* workspace, int64_t workspace_bytes, cyb_stream_t stream);
/* `kernels`: bit 0 = launch the streaming kernel (some block has 1 <= M <= cyb_grouped_gemm_skinny_rows() rows), bit 1 =
 * launch the DMMA kernel (some block has more); 3 launches both. */
int64_t cyb_grouped_gemm_skinny_rows(void);
int cyb_grouped_gemm_run(const void* a, const void* b, void* c, int dtype, const int64_t* out_mn,
                         const int64_t* out_off, const int64_t* out_pair_ptr, const int64_t* pairs,
                         int64_t nblocks, void* workspace, int kernels, cyb_stream_t stream);

/* ---- (a7,a9,a10) per-sector factorisations ----------------------------------------------- *
 * Batched over `nmat` independent row-major matrices inside arenas (device pointers a,u,s,...).
 * `desc_host` is a HOST table int64[nmat][6] = (m, n, in_off, out0_off, out1_off, out2_off) in
 * elements of the respective arena; the library sizes and owns its scratch memory. */

/* Thin SVD  A_i (m x n) = U_i diag(S_i) Vh_i, k = min(m, n), S descending;
 * replaces block_backend->matrix_svd per sector (abelian.cpp:3517, numpy.cpp:1247-1297).
 *  desc row = (m, n, a_off, u_off, s_off, vh_off); U_i is m x k, Vh_i is k x n (row-major),
 *  S real (float64 arena `s`).  info[nmat] (device): Jacobi sweeps used, negative = not converged. */
int cyb_batched_svd(const void* a, void* u, double* s, void* vh, const int64_t* desc_host,
                    int64_t nmat, int dtype, int64_t* info, cyb_stream_t stream);

/* Thin QR  A_i (m x n) = Q_i (m x k) R_i (k x n); replaces block_backend->matrix_qr(block,false)
 * (abelian.cpp:3123, numpy.cpp:1236-1245 scipy.linalg.qr economic).
 *  desc row = (m, n, a_off, q_off, r_off, unused). */
int cyb_batched_qr(const void* a, void* q, void* r, const int64_t* desc_host, int64_t nmat,
                   int dtype, cyb_stream_t stream);

/* Hermitian eigendecomposition A_i (n x n) = V_i diag(W_i) V_i^H, W ascending;
 * replaces block_backend->eigh (abelian.cpp:1781, numpy.cpp:658-680 np.linalg.eigh).
 *  desc row = (n, n, a_off, unused, w_off, v_off). */
int cyb_batched_eigh(const void* a, double* w, void* v, const int64_t* desc_host, int64_t nmat,
                     int dtype, int64_t* info, cyb_stream_t stream);

/* Writes leading-identity blocks: block i is (m_i x n_i) row-major with ones on the main
 * diagonal; replaces eye_matrix + slice for absent sectors (abelian.cpp:3524-3538) and
 * eye_data (abelian.cpp:1790-1826). */
int cyb_fill_eye(void* dst, const int64_t* dims, const int64_t* off, int64_t nmat, int dtype,
                 cyb_stream_t stream);

/* ---- (a8) truncation: tensor_backend.cpp:138-242 ----------------------------------------- *
 * Global cut over the concatenated singular values s[n] (any order, as
 * diagonal_tensor_to_block lays them out). Writes keep[n] (uint8 0/1, original order),
 * result[0] = err (sum of discarded s^2), result[1] = new_norm (sum of kept s^2),
 * result[2] = number kept (as double).
 * chi_max < 0 means "None"; svd_min < 0 means "None". */
int cyb_truncate_select(const double* s, int64_t n, int64_t chi_max, int64_t chi_min,
                        double degeneracy_tol, double trunc_cut, double svd_min,
                        int minimize_error, uint8_t* keep, double* result, cyb_stream_t stream);

/* cyb_truncate_select without any allocation at call time (no cudaMalloc / cudaMallocAsync / cudaFree): the scratch is a
 * file-scope static __device__ array of the library, sized for CYB_TRUNCATE_STATIC_MAX_N values (2.1 MiB, held by every GPU
 * that loads the library until the process exits).  For callers that cannot allocate: code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.  Same arguments and results as cyb_truncate_select;
 * n > CYB_TRUNCATE_STATIC_MAX_N is rejected with CYB_ERR_LIMIT.  All calls share the one array (it keeps the last call's
 * contents, none of which is read before it is rewritten): calls that overlap on the GPU -- different streams, or
 * different host threads without ordering -- race and must be serialised by the caller.
 * This entry point has no counterpart in the reference. */
#define CYB_TRUNCATE_STATIC_MAX_N 65536
int cyb_truncate_select_static(const double* s, int64_t n, int64_t chi_max, int64_t chi_min,
                               double degeneracy_tol, double trunc_cut, double svd_min,
                               int minimize_error, uint8_t* keep, double* result, cyb_stream_t stream);

/* ---- (f1) block-batched BLAS-1 used between the hot ops ---------------------------------- */
/* out[0] = sum |x|^2 over `n` elements (norm squared; abelian.cpp:2780-2791). */
int cyb_sumsq(const void* x, int64_t n, int dtype, double* out, cyb_stream_t stream);
/* y = alpha * x (+ beta * y) elementwise on arenas of n elements (real alpha/beta). */
int cyb_axpby(const void* x, void* y, int64_t n, int dtype, double alpha, double beta,
              cyb_stream_t stream);
/* y = 1 / x where |x| >= cutoff, else 0, on n real elements: the pseudo-inverse of the Schmidt values of a sweep
 * (BlockBackend::cutoff_inverse, src/block_backend/numpy.cpp:645-656, behind cyten.pinv(DiagonalTensor),
 * src/tensors/ops_algebra.cpp:1514-1515; toycodes/tenpy_toycodes/d_dmrg.py:228, 257). */
int cyb_cutoff_inverse(const double* x, double* y, int64_t n, double cutoff, cyb_stream_t stream);
/* scales row r of matrix i (m_i x n_i) by d[d_off[i] + r] (axis=0) or column (axis=1);
 * scale_axis / U*S products (abelian.cpp:3177-3232). `inv` != 0 divides. */
int cyb_scale_rows_cols(void* x, const int64_t* dims, const int64_t* off, const double* d,
                        const int64_t* d_off, int64_t nmat, int axis, int dtype,
                        cyb_stream_t stream);

/* ---- (f1) Krylov vector operations over whole arenas (one launch for all blocks) ------------ *
 * Segment tables (device int64): one row per pair of matching blocks, offsets in ELEMENTS of the operand
 * arenas (-1 = the operand has no such block) followed by the block size; `chunk0[nseg + 1]` = running
 * count of chunks of cyb_blas1_chunk_elems(dtype) elements per segment (ceil), `nchunk` = chunk0[nseg].
 * `work` = device scratch of cyb_blas1_workspace_bytes(), zero-initialised ONCE by the caller (the kernels
 * leave it re-armed); reductions are deterministic (fixed partial order, no floating-point atomics). */
int64_t cyb_blas1_workspace_bytes(void);
int64_t cyb_blas1_chunk_elems(int dtype);
/* out[0] + i out[1] = sum_seg sum_e conj?(x[off_x + e]) * y[off_y + e]; seg rows = (off_x, off_y, size).
 * conj_x bit 0: conjugate x; bit 1: write the real part only (out[1] = 0).
 * inner (abelian.cpp:2159-2211) and, with x == y, norm^2 (abelian.cpp:2780-2791). */
int cyb_block_dot(const void* x, const void* y, const int64_t* seg, const int64_t* chunk0, int64_t nseg,
                  int64_t nchunk, int dtype, int conj_x, double* out, void* work, cyb_stream_t stream);
/* out = a x + b y + c z per segment; seg rows = (off_x, off_y, off_z, off_out, size); x/y/z may be NULL.
 * coef = host double[6] (re, im of a, b, c); coef_dev = NULL or host array of 3 device pointers (NULL entries
 * allowed) to (re, im) factors multiplied onto a, b, c -- a Lanczos step never reads alpha back to the host.
 * sumsq != NULL: also writes ||out||^2 (fused norm).  `out` may alias an operand when the offsets coincide.
 * linear_combination / mul (abelian.cpp:2253-2301, 2759-2778), the axpy's of krylov_based.cpp:856-871. */
int cyb_block_lincomb(const void* x, const void* y, const void* z, void* out, const int64_t* seg,
                      const int64_t* chunk0, int64_t nseg, int64_t nchunk, int dtype, const double* coef,
                      const double* const* coef_dev, double* sumsq, void* work, cyb_stream_t stream);
/* scale_axis (abelian.cpp:3177-3232): out block = in block, viewed as (outer, dleg, inner), times
 * diag[d_off + r] along the middle axis.  desc (device) int64[nblk][6] = (off_in, off_out, outer, dleg,
 * inner, d_off); work0[nblk + 1] = running count of chunks of cyb_blas1_chunk_elems(dtype) elements per block
 * (ceil), nwork = work0[nblk]. */
int cyb_scale_axis(const void* in, void* out, const int64_t* desc, const int64_t* work0, int64_t nblk,
                   int64_t nwork, const void* diag, int dtype, int diag_dtype, cyb_stream_t stream);

/* ---- (f3) fused application of a small operator to the middle legs of a large tensor ------------------------- *
 * out[o][a][b] = sum_i C_g[o][i] * in[i][a][b] for every group g, a < na, b < nb: ONE streaming pass that replaces the
 * compose + permute_legs pairs of the two MPO applications inside the DMRG effective Hamiltonian (HEffective.matvec,
 * toycodes/tenpy_toycodes/d_dmrg.py:70-88; PlanarHEffective :94-116 -> src/tensors/planar.cpp).
 * groups (device) int64[ngroups][6] = (na, nb, nin, nout, desc_begin, coef_begin); desc (device) int64[.][2] =
 * (element offset, element stride per a) of the nin input rows, then of the nout output rows of each group;
 * coef (device): C_g row-major [nout][nin] elements of `dtype` at coef_begin; work0 (device) int64[ngroups + 1]: running
 * count of work units, na * ceil(nb / cyb_middle_apply_unit_cols()) per group; max_rows = max over groups of nin, nout
 * (<= cyb_middle_apply_max_rows(); <= 8 selects the kernel with half the registers). */
int64_t cyb_middle_apply_max_rows(void);
int64_t cyb_middle_apply_unit_cols(void);
int cyb_middle_apply(const void* in, void* out, const int64_t* groups, const int64_t* desc, const void* coef,
                     const int64_t* work0, int64_t ngroups, int64_t nwork, int64_t max_rows, int dtype, cyb_stream_t stream);

/* Keep-mask compaction (apply_mask / _mask_contract, abelian.cpp:2483-2582, 645-673):
 * out_i = in_i[idx, :] (axis 0) or in_i[:, idx] (axis 1).  desc (device) int64[nmat][6] =
 * (m_in, n_in, src_off, dst_off, idx_begin, idx_count); idx (device) concatenated index lists. */
int cyb_gather_rows_cols(const void* src, void* dst, const int64_t* desc, const int64_t* idx,
                         int64_t nmat, int axis, int dtype, cyb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CYTEN_B200_H */
