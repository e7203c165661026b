// One-sided BLOCK Jacobi for the large sectors of svd / eigh (included by factorize.cu); real and complex.
// complex128 sectors run on the SAME real DMMA machinery over the interleaved (re, im) view of the rows:
//   Re(R R^H) is the plain Gram matrix of the 2c-long real rows, Im(R R^H) the Gram product with the
//   B fragment taken from the neighbouring column with a sign (b'[2k] = -b[2k+1], b'[2k+1] = b[2k]);
//   Q R = Qr R + Qi R' with the same swizzle on the column index.  So a complex sector costs 2 DMMAs
//   per fragment pair on twice as many columns (the 4 real products of a complex product).
//
// The scalar kernel in factorize.cu streams two rows of W and Z through L2 for every plane rotation;
// for a 1000 x 1000 sector that is ~30 GB of traffic per sweep and no tensor-core work.  Here the rows
// are grouped in blocks of BJ_B = 32.  A work item is a pair of row blocks (I, J) of one matrix (or a
// single block, for the rotations inside a block); one CTA
//   1. forms the 64 x 64 Gram matrix G = R R^T of the item's rows R = [W_I; W_J] on the FP64 tensor
//      pipe (DMMA m8n8k4, operands staged global -> shared with cp.async, double buffered),
//   2. checks the item's pairs against the Jacobi criterion |g_pq| <= tol sqrt(g_pp g_qq) on this FRESH
//      Gram matrix; if some pair fails, runs one cyclic pass of plane rotations over the item's pairs
//      entirely in shared memory (32 disjoint rotations per step; G is updated two-sidedly, the product of
//      the rotations is accumulated in Q),
//   3. applies Q to the rows of W and (svd only: eigh reads its eigenvectors off W) of the accumulated transformation Z with DMMA GEMMs
//      ([W_I; W_J] <- Q [W_I; W_J]), in place, 64 columns at a time.
// The items of one round-robin round touch disjoint rows, so a round is one kernel launch over all
// large sectors of the batch; a sweep = 1 launch for the within-block pairs + (#blocks - 1) launches
// for the block pairs = every pair of rows exactly once (cyclic Jacobi ordering).  Convergence is decided
// only on fresh Gram matrices, never on the two-sidedly updated ones.
#pragma once

namespace cyb {

constexpr int BJ_B = 32;                 // rows per block
constexpr int BJ_R = 2 * BJ_B;           // rows per work item
constexpr int BJ_THREADS = 256;
constexpr int BJ_BK = 32;                // k-chunk of the Gram product
constexpr int BJ_LDT = BJ_BK + 4;        // 36: conflict-free fragment loads (ld = 4 mod 16)
constexpr int BJ_LDQ = BJ_R + 4;         // 68
constexpr int BJ_LDG = BJ_R + 1;         // 65
constexpr int BJ_BN = 64;                // column chunk of the update
constexpr int BJ_LDB = BJ_BN + 4;        // 68
constexpr int BJ_SQ = BJ_R * BJ_LDQ;                       // doubles
constexpr int BJ_SBUF_A = BJ_R * BJ_LDG + 2 * BJ_R * BJ_LDT;   // phases 1-2: G + two Gram tiles
constexpr int BJ_SBUF_B = 2 * BJ_R * BJ_LDB;                   // phase 3: two update tiles (aliases the above)
constexpr int BJ_SBUF = BJ_SBUF_A > BJ_SBUF_B ? BJ_SBUF_A : BJ_SBUF_B;
constexpr size_t BJ_SMEM = sizeof(double) * (size_t)(BJ_SQ + BJ_SBUF);
#ifndef CYB_BJ_KS
#define CYB_BJ_KS 128
#endif
#ifndef CYB_BJ_AC
#define CYB_BJ_AC 128
#endif
constexpr int BJ_KS = CYB_BJ_KS;         // columns per partial Gram matrix
constexpr int BJ_AC = CYB_BJ_AC;         // columns per apply work unit
constexpr int BJ_GST = BJ_KS / BJ_BK;        // k-chunks of a Gram slice: all of them are in flight at once
constexpr size_t BJ_SMEM_GRAM = sizeof(double) * (size_t)(BJ_GST * BJ_R * BJ_LDT);
constexpr size_t BJ_SMEM_ROT = sizeof(double) * (size_t)(2 * BJ_R * BJ_LDG);
constexpr size_t BJ_SMEM_APPLY = sizeof(double) * (size_t)(BJ_SQ + BJ_SBUF_B);
constexpr size_t BJ_SMEM_ROT_CX = 2 * BJ_SMEM_ROT;                                            // + Im G, Im Q
constexpr size_t BJ_SMEM_APPLY_CX = sizeof(double) * (size_t)(2 * BJ_SQ + BJ_SBUF_B);        // + Im Q

struct BjItem {
    int32_t mat;     // index into the FzDesc table
    int32_t i0, ni;  // rows [i0, i0 + ni) of block I  -> slots 0 .. 31
    int32_t j0, nj;  // rows [j0, j0 + nj) of block J  -> slots 32 .. 63 (nj == 0: within-block item)
    int32_t rowbase; // start of this matrix in the rowmap / norm2 arrays
    int32_t w0, nw;  // this item's partial Gram matrices: work units [w0, w0 + nw) of its round
};
// A round runs as three launches so that the whole GPU works on it (an item alone on one SM would be a
// ~150 us critical path per round, 31 rounds per sweep):
//   bj_gram_kernel   one CTA per (item, slice of BJ_KS columns): partial 64 x 64 Gram matrix -> Gpart
//   bj_rotate_kernel one CTA per item: sums the partials in a fixed order (deterministic), convergence
//                    test, one cyclic pass of rotations in shared memory -> Q (or skip flag)
//   bj_apply_kernel  one CTA per (item, W or Z, chunk of BJ_AC columns): rows <- Q rows, in place
struct BjWork {
    int32_t item;    // index into the round's item list
    int32_t arg;     // gram: slice index; apply: (which << 24) | chunk, which = 0 (W) / 1 (Z)
};
// i0 / j0 are POSITIONS in the norm-sorted order of the rows (de Rijk): position -> physical row through
// rowmap.  Sorting the rows by decreasing norm before every sweep gathers the (numerical) null space of a
// rank-deficient sector in the trailing blocks, which restores fast convergence (numpy experiment:
// 34 -> 12 sweeps for a rank-n/2 matrix); pairs of rows that are BOTH below the noise floor
// (eps * r * ||A||_F) are left alone until the significant part has converged and are orthogonalised in
// a second, much smaller, pass over the trailing blocks only.

__device__ __forceinline__ void bj_cp8(void* smem, const void* g, bool pred)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(s), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void bj_cp16(void* smem, const void* g, bool pred)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void bj_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void bj_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
// at most n (0 .. 3) of the most recently committed groups still pending; n folds to a constant in unrolled loops
__device__ __forceinline__ void bj_wait_pending(int n)
{
    switch (n) {
    case 3: bj_wait<3>(); break;
    case 2: bj_wait<2>(); break;
    case 1: bj_wait<1>(); break;
    default: bj_wait<0>(); break;
    }
}
static_assert(CYB_BJ_KS / 32 >= 1 && CYB_BJ_KS / 32 <= 4, "bj_gram_kernel keeps every 32-column chunk of a slice in flight (at most 4)");

// Programmatic dependent launch (CYB_BJ_PDL=1; off by default, see factorize.cu): the three kernels of a round are then
// launched with cudaLaunchAttributeProgrammaticStreamSerialization, so the CTAs of kernel N + 1 are scheduled (parameters
// fetched, prologue run) while kernel N still works; this wait returns once kernel N has completed and its writes are
// visible.  The trigger comes AFTER the wait: at most two kernels of a stream are in flight, and completion stays
// transitive (N + 1 cannot complete before N has).  Without the launch attribute both instructions are no-ops.
__device__ __forceinline__ void bj_pdl_enter()
{
    asm volatile("griddepcontrol.wait;\n" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
}

__device__ __forceinline__ void bj_dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// position of slot t in the sorted order (or -1)
__device__ __forceinline__ int bj_pos(const BjItem& it, int t)
{
    if (t < BJ_B) return t < it.ni ? it.i0 + t : -1;
    const int u = t - BJ_B;
    return u < it.nj ? it.j0 + u : -1;
}

// stage rows x [col0, col0 + NC) of a row-major matrix (leading dimension ld, ncols valid columns) into
// a shared tile [BJ_R][LDS]; missing rows / columns are zero-filled
template <int NC, int LDS>
__device__ __forceinline__ void bj_load_tile(double* tile, const double* __restrict__ base, int64_t ld,
                                             int64_t ncols, int64_t col0, const int* s_rows, bool vec)
{
    const int tid = threadIdx.x;
    if (vec) {
        constexpr int CH = NC / 2;
        for (int c = tid; c < BJ_R * CH; c += BJ_THREADS) {
            const int t = c / CH, k = (c % CH) * 2;
            const int row = s_rows[t];
            const bool ok = row >= 0 && col0 + k < ncols;    // ld even and ncols == ld or ncols even: pairs never straddle
            bj_cp16(tile + t * LDS + k, ok ? base + (int64_t)row * ld + col0 + k : base, ok);
        }
    } else {
        for (int c = tid; c < BJ_R * NC; c += BJ_THREADS) {
            const int t = c / NC, k = c % NC;
            const int row = s_rows[t];
            const bool ok = row >= 0 && col0 + k < ncols;
            bj_cp8(tile + t * LDS + k, ok ? base + (int64_t)row * ld + col0 + k : base, ok);
        }
    }
}

// [rows of the item] <- Q [rows of the item] for columns [cbeg, cend) of one row-major matrix, 64 columns
// per step, in place
template <bool CX>
__device__ __forceinline__ void bj_apply(const double* sQ, const double* sQi, double* sBuf, double* __restrict__ base, int64_t ld,
                                         int64_t ncols, const int* s_rows, int64_t cbeg, int64_t cend, bool preloaded = false)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const int wm0 = (warp >> 1) * 16, wn0 = (warp & 1) * 32;
    const bool vec = (ld & 1) == 0 && (ncols & 1) == 0;
    if (cend > ncols) cend = ncols;
    const int64_t nchunks = (cend - cbeg + BJ_BN - 1) / BJ_BN;
    if (!preloaded) {                        // (the caller may have requested the first tile already: bj_apply_kernel)
        bj_load_tile<BJ_BN, BJ_LDB>(sBuf, base, ld, cend, cbeg, s_rows, vec);
        bj_commit();
    }
    for (int64_t ch = 0; ch < nchunks; ++ch) {
        double* cur = sBuf + (ch & 1) * (BJ_R * BJ_LDB);
        if (ch + 1 < nchunks) bj_load_tile<BJ_BN, BJ_LDB>(sBuf + ((ch + 1) & 1) * (BJ_R * BJ_LDB), base, ld, cend, cbeg + (ch + 1) * BJ_BN, s_rows, vec);
        bj_commit();
        bj_wait<1>();
        __syncthreads();
        double acc[2][4][2];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
#pragma unroll 4
        for (int kk = 0; kk < BJ_R; kk += 4) {
            double a[2], b[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) a[i] = sQ[(wm0 + i * 8 + lr) * BJ_LDQ + kk + lc];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = cur[(kk + lc) * BJ_LDB + wn0 + j * 8 + lr];
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) bj_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            if (CX) {
                // + Qi R': R'[k][2n] = -R[k][2n+1], R'[k][2n+1] = R[k][2n]  (column parity = parity of lr)
                double ai[2], bs[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) ai[i] = sQi[(wm0 + i * 8 + lr) * BJ_LDQ + kk + lc];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double v = cur[(kk + lc) * BJ_LDB + ((wn0 + j * 8 + lr) ^ 1)];
                    bs[j] = (lr & 1) ? v : -v;
                }
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) bj_dmma(acc[i][j][0], acc[i][j][1], ai[i], bs[j]);
            }
        }
        const int64_t n0 = cbeg + ch * BJ_BN;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int row = s_rows[wm0 + i * 8 + lr];
            if (row < 0) continue;
            double* out = base + (int64_t)row * ld + n0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = wn0 + j * 8 + lc * 2;
                if (vec) {
                    if (n0 + col < cend) __stcg(reinterpret_cast<double2*>(out + col), make_double2(acc[i][j][0], acc[i][j][1]));
                } else {
                    if (n0 + col < cend) __stcg(out + col, acc[i][j][0]);
                    if (n0 + col + 1 < cend) __stcg(out + col + 1, acc[i][j][1]);
                }
            }
        }
        __syncthreads();   // everybody is done with `cur` before it is refilled two chunks later
    }
    bj_wait<0>();
}

struct BjRot { int p, q; double cs, sn, si; };   // rotation [[cs, -s], [conj(s), cs]], s = sn + i si (si = 0 for real sectors)

__device__ __forceinline__ void bj_fill_rows(int* s_rows, const BjItem& it, const int* __restrict__ rowmap)
{
    if (threadIdx.x < BJ_R) {
        const int pos = bj_pos(it, threadIdx.x);
        s_rows[threadIdx.x] = pos >= 0 ? rowmap[it.rowbase + pos] : -1;
    }
}

// ---- 1. partial Gram matrices -----------------------------------------------------------------
template <bool CX>
__global__ void __launch_bounds__(BJ_THREADS, CX ? 2 : 3)
bj_gram_kernel(const double* __restrict__ ws, const FzDesc* __restrict__ descs, const BjItem* __restrict__ items,
               const BjWork* __restrict__ works, const int* __restrict__ active, const int* __restrict__ rowmap,
               double* __restrict__ gpart)
{
    extern __shared__ __align__(16) double bj_smem[];
    bj_pdl_enter();
    double* sT = bj_smem;                    // BJ_GST tiles [64][36]
    __shared__ int s_rows[BJ_R];
    const BjWork wk = works[blockIdx.x];
    const BjItem it = items[wk.item];
    if (!active[it.mat]) return;
    bj_fill_rows(s_rows, it, rowmap);
    const FzDesc D = descs[it.mat];
    constexpr int CM = CX ? 2 : 1;            // real columns per element
    const int64_t c = D.c * CM;
    const double* W = ws + D.w_off * CM;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lr = lane >> 2, lc = lane & 3;
    const bool vec = (c & 1) == 0;
    const int wm0 = (warp >> 2) * 32, wn0 = (warp & 3) * 16;
    double acci[CX ? 4 : 1][2][2];
#pragma unroll
    for (int i = 0; i < (CX ? 4 : 1); ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) { acci[i][j][0] = 0.0; acci[i][j][1] = 0.0; }
    const int64_t k_beg = (int64_t)wk.arg * BJ_KS;
    const int64_t k_end = k_beg + BJ_KS < c ? k_beg + BJ_KS : c;
    __syncthreads();
    double acc[4][2][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
    const int64_t nk = (k_end - k_beg + BJ_BK - 1) / BJ_BK;
    // the whole slice (<= BJ_GST chunks of 32 columns) is requested up front: one L2 round trip instead of one per chunk
#pragma unroll
    for (int kc = 0; kc < BJ_GST; ++kc) {
        if (kc < nk) bj_load_tile<BJ_BK, BJ_LDT>(sT + kc * (BJ_R * BJ_LDT), W, c, k_end, k_beg + kc * BJ_BK, s_rows, vec);
        bj_commit();
    }
#pragma unroll
    for (int kc = 0; kc < BJ_GST; ++kc) {
        if (kc >= nk) break;
        const double* cur = sT + kc * (BJ_R * BJ_LDT);
        bj_wait_pending(BJ_GST - 1 - kc);    // chunks 0 .. kc have landed (the groups complete in order)
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BJ_BK; kk += 4) {
            double a[4], b[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = cur[(wm0 + i * 8 + lr) * BJ_LDT + kk + lc];
#pragma unroll
            for (int j = 0; j < 2; ++j) b[j] = cur[(wn0 + j * 8 + lr) * BJ_LDT + kk + lc];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) bj_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            if (CX) {
                // Im G[p][q] = sum_k (Wp[2k+1] Wq[2k] - Wp[2k] Wq[2k+1]): B fragment from the neighbouring column
                double bs[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const double v = cur[(wn0 + j * 8 + lr) * BJ_LDT + kk + (lc ^ 1)];
                    bs[j] = (lc & 1) ? v : -v;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 2; ++j) bj_dmma(acci[CX ? i : 0][j][0], acci[CX ? i : 0][j][1], a[i], bs[j]);
            }
        }
    }
    bj_wait<0>();
    double* out = gpart + (size_t)blockIdx.x * (BJ_R * BJ_R * CM);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int row = wm0 + i * 8 + lr, col = wn0 + j * 8 + lc * 2;
            __stcg(reinterpret_cast<double2*>(out + row * BJ_R + col), make_double2(acc[i][j][0], acc[i][j][1]));
            if (CX) __stcg(reinterpret_cast<double2*>(out + BJ_R * BJ_R + row * BJ_R + col), make_double2(acci[CX ? i : 0][j][0], acci[CX ? i : 0][j][1]));
        }
}

// ---- 2. convergence test + one cyclic pass of rotations ----------------------------------------------
constexpr int BJ_RT = 1024;              // threads of the rotation kernel

template <bool CX>
__global__ void __launch_bounds__(BJ_RT)
bj_rotate_kernel(const FzDesc* __restrict__ descs, const BjItem* __restrict__ items, const int* __restrict__ active,
                 int* __restrict__ flags, const double* __restrict__ floor2, int use_floor,
                 const double* __restrict__ gpart, double* __restrict__ qbuf, int* __restrict__ skip)
{
    extern __shared__ __align__(16) double bj_smem[];
    bj_pdl_enter();
    constexpr int CM = CX ? 2 : 1;
    constexpr int PL = BJ_R * BJ_LDG;        // one [64][65] plane
    double* sG = bj_smem;                    // Re G
    double* sQ = bj_smem + PL;               // Re Q
    double* sGi = bj_smem + 2 * PL;          // Im G, Im Q (complex sectors only)
    double* sQi = bj_smem + 3 * PL;
    __shared__ BjRot s_rot[BJ_B];
    __shared__ int s_any;
    const BjItem it = items[blockIdx.x];
    const int tid = threadIdx.x;
    if (!active[it.mat]) { if (tid == 0) skip[blockIdx.x] = 1; return; }
    const double fl = use_floor ? floor2[it.mat] : 0.0;
    const double tol = FZ_EPS * sqrt((double)descs[it.mat].c);
    const double tol2 = tol * tol;
    const bool cross = it.nj > 0;
    // G = sum of the partials, in slice order
    for (int e = tid; e < BJ_R * BJ_R; e += BJ_RT) {
        double g = 0.0, gi = 0.0;
        // fixed summation order; eight (then four) partials are in flight at a time (one L2 round trip instead of one each)
        int w = 0;
        for (; w + 7 < it.nw; w += 8) {      // real sectors of the chi = 4096 configuration: nw = 8 (974 columns / 128)
            const double* part = gpart + (size_t)(it.w0 + w) * (BJ_R * BJ_R * CM);
            double a[8], b[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                a[k] = __ldcg(part + (size_t)k * (BJ_R * BJ_R * CM) + e);
                if (CX) b[k] = __ldcg(part + (size_t)k * (BJ_R * BJ_R * CM) + BJ_R * BJ_R + e);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) { g += a[k]; if (CX) gi += b[k]; }
        }
        for (; w + 3 < it.nw; w += 4) {
            const double* part = gpart + (size_t)(it.w0 + w) * (BJ_R * BJ_R * CM);
            const double a0 = __ldcg(part + e), a1 = __ldcg(part + (size_t)(BJ_R * BJ_R * CM) + e);
            const double a2 = __ldcg(part + (size_t)2 * (BJ_R * BJ_R * CM) + e), a3 = __ldcg(part + (size_t)3 * (BJ_R * BJ_R * CM) + e);
            g += a0; g += a1; g += a2; g += a3;
            if (CX) {
                const double b0 = __ldcg(part + BJ_R * BJ_R + e), b1 = __ldcg(part + (size_t)(BJ_R * BJ_R * CM) + BJ_R * BJ_R + e);
                const double b2 = __ldcg(part + (size_t)2 * (BJ_R * BJ_R * CM) + BJ_R * BJ_R + e);
                const double b3 = __ldcg(part + (size_t)3 * (BJ_R * BJ_R * CM) + BJ_R * BJ_R + e);
                gi += b0; gi += b1; gi += b2; gi += b3;
            }
        }
        for (; w < it.nw; ++w) {
            const double* part = gpart + (size_t)(it.w0 + w) * (BJ_R * BJ_R * CM);
            g += __ldcg(part + e);
            if (CX) gi += __ldcg(part + BJ_R * BJ_R + e);
        }
        sG[(e / BJ_R) * BJ_LDG + e % BJ_R] = g;
        if (CX) sGi[(e / BJ_R) * BJ_LDG + e % BJ_R] = gi;
    }
    if (tid == 0) s_any = 0;
    __syncthreads();
    {
        int bad = 0;
        for (int e = tid; e < BJ_B * BJ_B; e += BJ_RT) {
            int p, q;
            bool valid;
            if (cross) { p = e / BJ_B; q = BJ_B + e % BJ_B; valid = p < it.ni && q - BJ_B < it.nj; }
            else { p = e / BJ_B; q = e % BJ_B; valid = p < q && q < it.ni; }
            if (valid) {
                const double al = sG[p * BJ_LDG + p], be = sG[q * BJ_LDG + q], ga = sG[p * BJ_LDG + q];
                const double gi = CX ? sGi[p * BJ_LDG + q] : 0.0;
                bad |= (al >= FZ_TINY && be >= FZ_TINY && !(al <= fl && be <= fl) && ga * ga + gi * gi > tol2 * al * be);
            }
        }
        if (bad) s_any = 1;
        __syncthreads();
        if (!s_any) { if (tid == 0) skip[blockIdx.x] = 1; return; }   // every pair of this item is orthogonal
        if (tid == 0) { flags[it.mat] = 1; skip[blockIdx.x] = 0; }
    }
    for (int e = tid; e < BJ_R * BJ_R; e += BJ_RT) {
        sQ[(e / BJ_R) * BJ_LDG + e % BJ_R] = (e / BJ_R == e % BJ_R) ? 1.0 : 0.0;
        if (CX) sQi[(e / BJ_R) * BJ_LDG + e % BJ_R] = 0.0;
    }
    const int nsteps = cross ? BJ_B : BJ_B - 1;
    const int nrot = cross ? BJ_B : BJ_B / 2;
    __syncthreads();
    if (cross) {
        // (What bounds a step, profiles/r2/svd_groups_r2f.txt: not the FP64 pipe -- tools/dfma_probe.cu measures 61 lane-FMA /
        //  clk / SM, and computing only the upper triangle of G changed nothing -- but the shared-memory pipe, which serves
        //  the ~1000 warp-wide 8-byte requests of a step in order, with warp 0's dependent loads queued among them.)
        // Block pairs (every round but the first of a sweep): ONE barrier per step, rotations computed once.
        // Rotation b of step s pairs row p_b = b with row q_b = 32 + (b + s) % 32.  Thread (warp a, lane b) owns the
        // 2 x 2 block G[{p_a, q_a}][{p_b, q_b}] and columns a, a + 32 of the two rows of Q that rotation b mixes.  The
        // rotations of step s + 1 need the new diagonal and the new pivots G[p_b][q_{b+1}]: exactly the blocks (b, b) and
        // (b, b + 1).  Warp 0 updates those 64 blocks FIRST (two per lane), derives the next rotations from its own
        // registers (+ one shuffle) and publishes them in the other half of a double-buffered table while the other
        // warps are still updating their blocks -- so the serial sqrt / divide / rsqrt chain of a step hides behind the
        // block updates instead of sitting between two barriers.
        __shared__ double s_cs[2][BJ_B], s_sn[2][BJ_B], s_si[2][BJ_B];
        const int wa = tid >> 5, lb = tid & 31;
        auto rotation = [&](int pb, int qb, double al, double be, double gr, double gi, double& cs, double& sn, double& si) {
            cs = 1.0; sn = 0.0; si = 0.0;
            if (pb < it.ni && qb - BJ_B < it.nj) {
                const double g2 = gr * gr + gi * gi;
                if (al >= FZ_TINY && be >= FZ_TINY && !(al <= fl && be <= fl) && g2 > tol2 * al * be) {
                    const double rg = CX ? rsqrt(g2) : 0.0;
                    const double gabs = CX ? g2 * rg : fabs(gr);
                    // with h = sqrt(d^2 + g^2), g = 2 |gamma|:  cos^2 = (h + |d|) / (2 h),  sin = sign(d) g / (2 h cos) -- the
                    // same rotation as t = sign(d) g / (|d| + h), cos = 1 / sqrt(1 + t^2), with two rsqrt in series
                    // (75 cycles each) instead of sqrt + divide + rsqrt (100 + 80 + 75) on the chain of a step
                    const double d = be - al, g2x = 2.0 * gabs;
                    const double rh = rsqrt(d * d + g2x * g2x);
                    const double c2 = 0.5 + 0.5 * (fabs(d) * rh);
                    const double rc = rsqrt(c2);
                    cs = c2 * rc;
                    double s0 = (0.5 * g2x) * rh * rc;
                    if (d < 0.0) s0 = -s0;
                    if (CX) { sn = s0 * gr * rg; si = s0 * gi * rg; }
                    else sn = gr >= 0.0 ? s0 : -s0;
                }
            }
        };
        // one 2 x 2 block: rows rotated by (csa, sna, sia), columns by (csb, snb, sib); returns the new (0,0), (0,1), (1,1)
        auto block = [&](int pa, int qa, int pb, int qb, double csa, double sna, double sia, double csb, double snb, double sib,
                         double& n00, double& n01, double& n01i, double& n11) {
            const int i00 = pa * BJ_LDG + pb, i01 = pa * BJ_LDG + qb, i10 = qa * BJ_LDG + pb, i11 = qa * BJ_LDG + qb;
            if (!CX) {
                const double g00 = sG[i00], g01 = sG[i01], g10 = sG[i10], g11 = sG[i11];
                const double h00 = csa * g00 - sna * g10, h01 = csa * g01 - sna * g11;
                const double h10 = sna * g00 + csa * g10, h11 = sna * g01 + csa * g11;
                n00 = csb * h00 - snb * h01;
                n01 = snb * h00 + csb * h01;
                const double n10 = csb * h10 - snb * h11;
                n11 = snb * h10 + csb * h11;
                n01i = 0.0;
                sG[i00] = n00; sG[i01] = n01; sG[i10] = n10; sG[i11] = n11;
            } else {
                const double g00r = sG[i00], g01r = sG[i01], g10r = sG[i10], g11r = sG[i11];
                const double g00i = sGi[i00], g01i = sGi[i01], g10i = sGi[i10], g11i = sGi[i11];
                const double h00r = csa * g00r - (sna * g10r - sia * g10i), h00i = csa * g00i - (sna * g10i + sia * g10r);
                const double h01r = csa * g01r - (sna * g11r - sia * g11i), h01i = csa * g01i - (sna * g11i + sia * g11r);
                const double h10r = (sna * g00r + sia * g00i) + csa * g10r, h10i = (sna * g00i - sia * g00r) + csa * g10i;
                const double h11r = (sna * g01r + sia * g01i) + csa * g11r, h11i = (sna * g01i - sia * g01r) + csa * g11i;
                n00 = csb * h00r - (snb * h01r + sib * h01i);
                const double n00i = csb * h00i - (snb * h01i - sib * h01r);
                n01 = (snb * h00r - sib * h00i) + csb * h01r;
                n01i = (snb * h00i + sib * h00r) + csb * h01i;
                const double n10 = csb * h10r - (snb * h11r + sib * h11i), n10i = csb * h10i - (snb * h11i - sib * h11r);
                n11 = (snb * h10r - sib * h10i) + csb * h11r;
                const double n11i = (snb * h10i + sib * h10r) + csb * h11i;
                sG[i00] = n00; sGi[i00] = n00i; sG[i01] = n01; sGi[i01] = n01i;
                sG[i10] = n10; sGi[i10] = n10i; sG[i11] = n11; sGi[i11] = n11i;
            }
        };
        // Real sectors: Q lives in registers during the pass (shared-memory bandwidth bounds a step: 128 B / clk / SM against
        // 64 KB of G and, before, 64 KB of Q read and written per step; chi = 4096: 46.3 -> 42.2 us per launch).  Complex
        // sectors keep Q in shared memory: with 1024 threads the kernel has 64 registers per thread and the four extra
        // doubles spill.
        constexpr bool REGQ = !CX;
        double qp_r[2], qq_r[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            qp_r[h] = (wa + h * BJ_B == lb) ? 1.0 : 0.0;
            qq_r[h] = (wa + h * BJ_B == BJ_B + lb) ? 1.0 : 0.0;
        }
        if (wa == 0) {                      // rotations of step 0 straight from G
            double cs, sn, si;
            const int q0 = BJ_B + lb;
            rotation(lb, q0, sG[lb * BJ_LDG + lb], sG[q0 * BJ_LDG + q0], sG[lb * BJ_LDG + q0], CX ? sGi[lb * BJ_LDG + q0] : 0.0, cs, sn, si);
            s_cs[0][lb] = cs; s_sn[0][lb] = sn; s_si[0][lb] = si;
        }
        __syncthreads();
        for (int step = 0; step < BJ_B; ++step) {
            const int par = step & 1, np = par ^ 1;
            const int pb = lb, qb = BJ_B + ((lb + step) & (BJ_B - 1));
            const int pa = wa, qa = BJ_B + ((wa + step) & (BJ_B - 1));
            const double cs = s_cs[par][lb], sn = s_sn[par][lb], si = CX ? s_si[par][lb] : 0.0;
            double n00, n01, n01i, n11;
            if (wa == 0) {
                // the blocks (b, b) and (b, b + 1) first, then the rotations of the next step
                const int l1 = (lb + 1) & (BJ_B - 1);
                const int p1 = l1, q1 = BJ_B + ((l1 + step) & (BJ_B - 1));
                const double cs1 = s_cs[par][l1], sn1 = s_sn[par][l1], si1 = CX ? s_si[par][l1] : 0.0;
                double d00, dxx, dxi, d11;
                block(pb, qb, pb, qb, cs, sn, si, cs, sn, si, d00, dxx, dxi, d11);          // (b, b): new diagonal of p_b, q_b
                block(pb, qb, p1, q1, cs, sn, si, cs1, sn1, si1, n00, n01, n01i, n11);      // (b, b+1): n01 = new G[p_b][q_{b+1}]
                const double be = __shfl_sync(0xffffffffu, d11, l1);                        // new G[q_{b+1}][q_{b+1}]
                double c2, s2, i2;
                rotation(pb, q1, d00, be, n01, n01i, c2, s2, i2);
                s_cs[np][lb] = c2; s_sn[np][lb] = s2; s_si[np][lb] = i2;
            }
            const bool special = lb == wa || lb == ((wa + 1) & (BJ_B - 1));
            if (!special) {
                const double csa = s_cs[par][wa], sna = s_sn[par][wa], sia = CX ? s_si[par][wa] : 0.0;
                block(pa, qa, pb, qb, csa, sna, sia, cs, sn, si, n00, n01, n01i, n11);
            }
            if (REGQ) {
                // Q <- J Q in registers: rows p_b and q_b, columns wa and wa + 32; the p row stays with the lane, the q row
                // of the next step (q_b(s + 1) = q_{b+1}(s)) comes from lane b + 1
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const double x = qp_r[h], y = qq_r[h];
                    qp_r[h] = cs * x - sn * y;
                    qq_r[h] = __shfl_sync(0xffffffffu, sn * x + cs * y, (lb + 1) & (BJ_B - 1));
                }
            } else if (sn != 0.0 || si != 0.0) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int ip = pb * BJ_LDG + wa + h * BJ_B, iq = qb * BJ_LDG + wa + h * BJ_B;
                    const double xr = sQ[ip], xi = sQi[ip], yr = sQ[iq], yi = sQi[iq];
                    sQ[ip] = cs * xr - (sn * yr - si * yi);   sQi[ip] = cs * xi - (sn * yi + si * yr);
                    sQ[iq] = (sn * xr + si * xi) + cs * yr;   sQi[iq] = (sn * xi - si * xr) + cs * yi;
                }
            }
            __syncthreads();
        }
        if (REGQ) {                         // after 32 shifts the q row of lane b is row 32 + b again
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                sQ[lb * BJ_LDG + wa + h * BJ_B] = qp_r[h];
                sQ[(BJ_B + lb) * BJ_LDG + wa + h * BJ_B] = qq_r[h];
            }
            __syncthreads();
        }
    } else
    for (int step = 0; step < nsteps; ++step) {
        if (tid < nrot) {
            int p, q;
            bool valid;
            if (cross) {
                p = tid; q = BJ_B + (tid + step) % BJ_B;
                valid = p < it.ni && q - BJ_B < it.nj;
            } else {
                if (tid == 0) { p = BJ_B - 1; q = step; }
                else { p = (step + tid) % (BJ_B - 1); q = (step - tid + (BJ_B - 1)) % (BJ_B - 1); }
                if (p > q) { const int t = p; p = q; q = t; }
                valid = q < it.ni;
            }
            double cs = 1.0, sn = 0.0, si = 0.0;
            if (valid) {
                const double al = sG[p * BJ_LDG + p], be = sG[q * BJ_LDG + q], gr = sG[p * BJ_LDG + q];
                const double gi = CX ? sGi[p * BJ_LDG + q] : 0.0;
                const double g2 = gr * gr + gi * gi;
                if (al >= FZ_TINY && be >= FZ_TINY && !(al <= fl && be <= fl) && g2 > tol2 * al * be) {
                    // rows x_p, x_q with <x_p, x_q> = ga = |ga| e^{i phi}:  x_p' = cs x_p - s x_q,
                    // x_q' = conj(s) x_p + cs x_q,  s = sn e^{i phi}  (real sectors: phi = 0 or pi)
                    // t = sign(zeta) / (|zeta| + sqrt(1 + zeta^2)), zeta = (be - al) / (2 |ga|), written without
                    // the division by ga (one sqrt, one division, one rsqrt on the serial path of a step)
                    const double rg = CX ? rsqrt(g2) : 0.0;
                    const double gabs = CX ? g2 * rg : fabs(gr);
                    const double d = be - al, g2x = 2.0 * gabs;
                    const double t = (d >= 0.0 ? g2x : -g2x) / (fabs(d) + sqrt(d * d + g2x * g2x));
                    cs = rsqrt(1.0 + t * t);
                    const double s0 = cs * t;
                    if (CX) { sn = s0 * gr * rg; si = s0 * gi * rg; }
                    else sn = gr >= 0.0 ? s0 : -s0;
                }
            }
            s_rot[tid].p = p; s_rot[tid].q = q; s_rot[tid].cs = cs; s_rot[tid].sn = sn; s_rot[tid].si = si;
        }
        __syncthreads();
        // G <- J G J^H, one 2x2 block G[{p,q}][{p',q'}] per (rotation a, rotation b): the blocks are disjoint,
        // so rows and columns are updated in ONE phase;  Q <- J Q (rows p, q) in the same phase
        for (int e = tid; e < nrot * nrot; e += BJ_RT) {
            const BjRot Ra = s_rot[e / nrot], Rb = s_rot[e % nrot];
            if (Ra.sn == 0.0 && Rb.sn == 0.0 && Ra.si == 0.0 && Rb.si == 0.0) continue;
            const int i00 = Ra.p * BJ_LDG + Rb.p, i01 = Ra.p * BJ_LDG + Rb.q;
            const int i10 = Ra.q * BJ_LDG + Rb.p, i11 = Ra.q * BJ_LDG + Rb.q;
            if (!CX) {
                const double g00 = sG[i00], g01 = sG[i01], g10 = sG[i10], g11 = sG[i11];
                // rows:  [x; y] <- [cs x - sn y; sn x + cs y]
                const double h00 = Ra.cs * g00 - Ra.sn * g10, h01 = Ra.cs * g01 - Ra.sn * g11;
                const double h10 = Ra.sn * g00 + Ra.cs * g10, h11 = Ra.sn * g01 + Ra.cs * g11;
                // columns with rotation b
                sG[i00] = Rb.cs * h00 - Rb.sn * h01;
                sG[i01] = Rb.sn * h00 + Rb.cs * h01;
                sG[i10] = Rb.cs * h10 - Rb.sn * h11;
                sG[i11] = Rb.sn * h10 + Rb.cs * h11;
            } else {
                const double g00r = sG[i00], g01r = sG[i01], g10r = sG[i10], g11r = sG[i11];
                const double g00i = sGi[i00], g01i = sGi[i01], g10i = sGi[i10], g11i = sGi[i11];
                // rows: h0 = cs g0 - s g1,  h1 = conj(s) g0 + cs g1   (s = sn + i si of rotation a)
                const double h00r = Ra.cs * g00r - (Ra.sn * g10r - Ra.si * g10i), h00i = Ra.cs * g00i - (Ra.sn * g10i + Ra.si * g10r);
                const double h01r = Ra.cs * g01r - (Ra.sn * g11r - Ra.si * g11i), h01i = Ra.cs * g01i - (Ra.sn * g11i + Ra.si * g11r);
                const double h10r = (Ra.sn * g00r + Ra.si * g00i) + Ra.cs * g10r, h10i = (Ra.sn * g00i - Ra.si * g00r) + Ra.cs * g10i;
                const double h11r = (Ra.sn * g01r + Ra.si * g01i) + Ra.cs * g11r, h11i = (Ra.sn * g01i - Ra.si * g01r) + Ra.cs * g11i;
                // columns: new[.][p] = cs h_p - conj(s) h_q,  new[.][q] = s h_p + cs h_q   (rotation b)
                sG[i00] = Rb.cs * h00r - (Rb.sn * h01r + Rb.si * h01i);  sGi[i00] = Rb.cs * h00i - (Rb.sn * h01i - Rb.si * h01r);
                sG[i01] = (Rb.sn * h00r - Rb.si * h00i) + Rb.cs * h01r;  sGi[i01] = (Rb.sn * h00i + Rb.si * h00r) + Rb.cs * h01i;
                sG[i10] = Rb.cs * h10r - (Rb.sn * h11r + Rb.si * h11i);  sGi[i10] = Rb.cs * h10i - (Rb.sn * h11i - Rb.si * h11r);
                sG[i11] = (Rb.sn * h10r - Rb.si * h10i) + Rb.cs * h11r;  sGi[i11] = (Rb.sn * h10i + Rb.si * h10r) + Rb.cs * h11i;
            }
        }
        for (int e = tid; e < nrot * BJ_R; e += BJ_RT) {
            const BjRot R = s_rot[e / BJ_R];
            if (R.sn == 0.0 && R.si == 0.0) continue;
            const int col = e % BJ_R;
            const int ip = R.p * BJ_LDG + col, iq = R.q * BJ_LDG + col;
            if (!CX) {
                const double x = sQ[ip], y = sQ[iq];
                sQ[ip] = R.cs * x - R.sn * y;
                sQ[iq] = R.sn * x + R.cs * y;
            } else {
                const double xr = sQ[ip], xi = sQi[ip], yr = sQ[iq], yi = sQi[iq];
                sQ[ip] = R.cs * xr - (R.sn * yr - R.si * yi);   sQi[ip] = R.cs * xi - (R.sn * yi + R.si * yr);
                sQ[iq] = (R.sn * xr + R.si * xi) + R.cs * yr;   sQi[iq] = (R.sn * xi - R.si * xr) + R.cs * yi;
            }
        }
        __syncthreads();
    }
    double* out = qbuf + (size_t)blockIdx.x * (BJ_R * BJ_R * CM);
    for (int e = tid; e < BJ_R * BJ_R; e += BJ_RT) {
        __stcg(out + e, sQ[(e / BJ_R) * BJ_LDG + e % BJ_R]);
        if (CX) __stcg(out + BJ_R * BJ_R + e, sQi[(e / BJ_R) * BJ_LDG + e % BJ_R]);
    }
}

// ---- 3. rows <- Q rows -------------------------------------------------------------------------------
template <bool CX>
__global__ void __launch_bounds__(BJ_THREADS, CX ? 1 : 2)
bj_apply_kernel(double* __restrict__ ws, const FzDesc* __restrict__ descs, const BjItem* __restrict__ items,
                const BjWork* __restrict__ works, const int* __restrict__ rowmap, const double* __restrict__ qbuf,
                const int* __restrict__ skip)
{
    extern __shared__ __align__(16) double bj_smem[];
    bj_pdl_enter();
    constexpr int CM = CX ? 2 : 1;
    double* sQ = bj_smem;                              // Re Q [64][68]
    double* sQi = bj_smem + BJ_SQ;                     // Im Q (complex sectors only)
    double* sBuf = bj_smem + (CX ? 2 : 1) * BJ_SQ;     // two tiles [64][68]
    __shared__ int s_rows[BJ_R];
    const BjWork wk = works[blockIdx.x];
    if (skip[wk.item]) return;
    const BjItem it = items[wk.item];
    bj_fill_rows(s_rows, it, rowmap);
    // Q (32 KB, L2) and the first tile of the rows are requested together: Q into registers, then -- as soon as the row
    // table is visible -- the tile with cp.async, and only then Q goes to shared memory (two round trips used to be serial)
    const double* q = qbuf + (size_t)wk.item * (BJ_R * BJ_R * CM);
    constexpr int QV = BJ_R * BJ_R / 2 / BJ_THREADS;       // double2 per thread
    double2 qv[QV], qvi[CX ? QV : 1];
#pragma unroll
    for (int k = 0; k < QV; ++k) {
        qv[k] = __ldcg(reinterpret_cast<const double2*>(q) + threadIdx.x + k * BJ_THREADS);
        if (CX) qvi[k] = __ldcg(reinterpret_cast<const double2*>(q + BJ_R * BJ_R) + threadIdx.x + k * BJ_THREADS);
    }
    const FzDesc D = descs[it.mat];
    const int which = wk.arg >> 24;
    const int64_t cbeg = (int64_t)(wk.arg & 0xffffff) * BJ_AC;
    double* base = ws + (which == 0 ? D.w_off : D.z_off) * CM;
    const int64_t ld = (which == 0 ? D.c : D.r) * CM;
    __syncthreads();                                       // s_rows
    {
        const int64_t cend = cbeg + BJ_AC < ld ? cbeg + BJ_AC : ld;
        bj_load_tile<BJ_BN, BJ_LDB>(sBuf, base, ld, cend, cbeg, s_rows, (ld & 1) == 0);
        bj_commit();
    }
#pragma unroll
    for (int k = 0; k < QV; ++k) {
        const int e = threadIdx.x + k * BJ_THREADS;
        const int row = (2 * e) / BJ_R, col = (2 * e) % BJ_R;
        sQ[row * BJ_LDQ + col] = qv[k].x;
        sQ[row * BJ_LDQ + col + 1] = qv[k].y;
        if (CX) {
            sQi[row * BJ_LDQ + col] = qvi[k].x;
            sQi[row * BJ_LDQ + col + 1] = qvi[k].y;
        }
    }
    __syncthreads();
    bj_apply<CX>(sQ, sQi, sBuf, base, ld, ld, s_rows, cbeg, cbeg + BJ_AC, true);
}

// squared row norms of the working copies: one warp per row, grid-stride over (matrix, row)
__global__ void __launch_bounds__(256)
bj_norms_kernel(const double* __restrict__ ws, const FzDesc* __restrict__ descs, const int64_t* __restrict__ mats,
                const int* __restrict__ rowbase, int nmats, double* __restrict__ norm2, int cm)
{
    const int lane = threadIdx.x & 31;
    const int64_t gwarp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = rowbase[nmats];
    for (int64_t g = gwarp; g < total; g += nwarps) {
        int lo = 0, hi = nmats;
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (rowbase[mid] <= g) lo = mid; else hi = mid; }
        const FzDesc D = descs[mats[lo]];
        const int64_t cr = D.c * cm;     // real columns (cm = 2: interleaved complex rows)
        const double* x = ws + D.w_off * cm + (g - rowbase[lo]) * cr;
        double acc = 0.0;
        for (int64_t q = lane; q < cr; q += 32) { const double v = __ldcg(x + q); acc = fma(v, v, acc); }
        acc = warp_sum(acc);
        if (lane == 0) norm2[g] = acc;
    }
}

// one CTA per matrix: rows ranked by decreasing norm (ties: lower row first) -> rowmap; with `init`
// also the noise floor (eps r)^2 ||A||_F^2; nnoise = number of rows at or below the floor
__global__ void __launch_bounds__(1024)
bj_sort_kernel(const FzDesc* __restrict__ descs, const int64_t* __restrict__ mats, const int* __restrict__ rowbase,
               const double* __restrict__ norm2, int* __restrict__ rowmap, double* __restrict__ floor2,
               int* __restrict__ nnoise, const int* __restrict__ active, int init)
{
    const int64_t mat = mats[blockIdx.x];
    if (!init && !active[mat]) return;
    const int r = (int)descs[mat].r;
    const int base = rowbase[blockIdx.x];
    const double* key = norm2 + base;
    __shared__ double s_red[32];
    __shared__ int s_cnt;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (init) {
        double acc = 0.0;
        for (int i = tid; i < r; i += 1024) acc += key[i];
        acc = warp_sum(acc);
        if (lane == 0) s_red[warp] = acc;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < 32; ++w) t += s_red[w];
            const double er = FZ_EPS * (double)r;
            floor2[mat] = er * er * t;
        }
        __syncthreads();
    }
    const double fl = floor2[mat];
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < r; i += 1024) {
        const double ki = key[i];
        int rank = 0;
        for (int j = 0; j < r; ++j) {
            const double kj = key[j];
            rank += (kj > ki) || (kj == ki && j < i);
        }
        rowmap[base + rank] = i;
        cnt += (ki <= fl);
    }
    if (cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (tid == 0) nnoise[mat] = s_cnt;
}

}  // namespace cyb
