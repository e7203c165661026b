// (a6) grouped per-sector contraction GEMMs on the FP64 tensor pipe.
//
//   C_t (M_t x N_t) = sum_{p in pairs(t)} A_p (M_t x K_p) * B_p (K_p x N_t)      (all row-major)
//
// replaces the matrix_dot / operator+ loop of abelian_compose_worker
// (src/backends/abelian.cpp:1424-1450; NumpyBlockBackend::matrix_dot = np.dot, numpy.cpp:1219-1225).
//
// Blackwell has no f64 kind on tcgen05, so float64 / complex128 run on warp-level DMMA: on sm_100a
// every mma.sync f64 shape lowers to DMMA.8x8x4 (checked with cuobjdump), which is what is issued
// here directly.  One persistent CTA per SM pulls output tiles from a device work list (largest
// block first, atomic ticket), and for each tile walks the FLATTENED k-loop over all matched
// (A,B) block pairs of that output block: the sum over pairs is accumulated in registers, there is
// no operator+ pass and no intermediate block.  Operands are staged global->shared with a
// multi-stage cp.async pipeline (zero-filled at ragged edges; 16-byte copies when the leading
// dimension is even, 8-byte otherwise -- blocks have arbitrary odd extents, so TMA tensor maps do
// not apply), shared-memory rows are padded so every fragment load is bank-conflict free.
// complex128 = 4 real DMMAs per fragment pair on split re/im accumulators.
#include "common.cuh"
#include <cstdlib>

namespace cyb {

__device__ __forceinline__ void cp_async_8(void* smem, const void* g, bool pred)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(s), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_16(void* smem, const void* g, bool pred)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ double neg_bits(double x)
{
    return __longlong_as_double(__double_as_longlong(x) ^ (long long)0x8000000000000000ull);
}

// One operand tile global -> shared.  Thread t copies chunk column cc = t % CC (a chunk = CD doubles: 16 or 8 bytes) of the
// rows r_i = t / CC + i * (THREADS / CC): the same assignment as "chunk c = t + i * THREADS" when CC divides THREADS, but
// cc, the first row and the validity of the column are computed once, and the source pointer advances by a constant.
// `src` is the tile origin in global memory (row stride ld doubles), rows_valid / cols_valid (doubles) bound the block;
// chunks outside are zero-filled (cp.async src-size 0 from a valid dummy address).
template <int ROWS, int CC, int CD, int LDS, int THREADS>
__device__ __forceinline__ void load_tile(double* __restrict__ dst, const double* __restrict__ src, int64_t ld,
                                          int rows_valid, int cols_valid, const double* __restrict__ dummy, int tid)
{
    static_assert(THREADS % CC == 0 && ROWS % (THREADS / CC) == 0, "load_tile: chunk columns must divide the CTA");
    constexpr int RS = THREADS / CC, IT = ROWS / RS;
    const int cc = tid % CC, r0 = tid / CC;
    const bool colok = cc * CD < cols_valid;
    const double* p = src + (int64_t)r0 * ld + cc * CD;
    double* d = dst + r0 * LDS + cc * CD;
    const int64_t step = (int64_t)RS * ld;
#pragma unroll
    for (int i = 0; i < IT; ++i) {
        const bool ok = colok && (r0 + i * RS < rows_valid);
        if (CD == 2) cp_async_16(d + i * RS * LDS, ok ? p : dummy, ok);
        else cp_async_8(d + i * RS * LDS, ok ? p : dummy, ok);
        p += step;
    }
}

struct GemmTables {
    const int64_t* out_mn;        // [nblocks][2]
    const int64_t* out_off;       // [nblocks]
    const int64_t* out_pair_ptr;  // [nblocks + 1]
    const int64_t* pairs;         // [npairs][3] = (a_off, b_off, K)
    const int64_t* order;         // [nblocks] processing order (largest first)
    const int64_t* tile_start;    // [nblocks + 1] prefix of tile counts in `order`
    unsigned long long* ticket;   // dynamic tile counter
    int64_t nblocks;
};

// Output blocks with M <= SK_MAXM rows (an MPO tensor applied to a big operand: M, K of a few units, N
// up to 10^5) have nothing for an 8x8x4 tensor-core tile to chew on: they are HBM-bound and go to the
// streaming kernel below (units of SK_COLS columns); everything else goes to the DMMA kernel.
constexpr int SK_MAXM = 8;
constexpr int SK_THREADS = 256;
constexpr int SK_CHUNK = 2 * SK_THREADS;   // columns one CTA covers per pass
constexpr int SK_PASSES = 4;               // passes per work unit (amortises the unit lookup)
constexpr int SK_COLS = SK_CHUNK * SK_PASSES;

template <int BM, int BN, bool SKINNY>
__host__ __device__ __forceinline__ int64_t n_units(int64_t M, int64_t N)
{
    if (M <= 0 || N <= 0) return 0;
    if (SKINNY) return M <= SK_MAXM ? (N + SK_COLS - 1) / SK_COLS : 0;
    return M > SK_MAXM ? ((M + BM - 1) / BM) * ((N + BN - 1) / BN) : 0;
}

__host__ __device__ __forceinline__ int64_t sk_units(int64_t M, int64_t N, int64_t ucols)
{
    return (M > 0 && M <= SK_MAXM && N > 0) ? (N + ucols - 1) / ucols : 0;
}

// Unit width of a streaming block: a unit moves ucols * (J + M) doubles (J rows of B read, M rows of C
// written), so the width is chosen per block such that every unit carries about `target` doubles --
// units of equal cost, handed out by an atomic ticket, keep all CTAs busy until the end.
__host__ __device__ __forceinline__ int64_t sk_ucols(int64_t target, int64_t J, int64_t M)
{
    const int64_t rows = J + M > 0 ? J + M : 1;
    int64_t u = (target / rows + SK_CHUNK - 1) / SK_CHUNK * SK_CHUNK;
    if (u < SK_CHUNK) u = SK_CHUNK;
    if (u > 64 * SK_CHUNK) u = 64 * SK_CHUNK;
    return u;
}

// single CTA: cost-sorted processing order, tile prefix, ticket reset.
// SKINNY = false: tiles of the DMMA kernel (blocks with M > SK_MAXM); true: column chunks of the
// streaming kernel (blocks with M <= SK_MAXM).
template <int BM, int BN, bool SKINNY>
__global__ void __launch_bounds__(1024) gemm_schedule_kernel(const int64_t* __restrict__ out_mn,
                                                             const int64_t* __restrict__ out_pair_ptr,
                                                             const int64_t* __restrict__ pairs,
                                                             const int64_t* __restrict__ subset,
                                                             int64_t nblocks, double* __restrict__ cost,
                                                             int64_t* __restrict__ order,
                                                             int64_t* __restrict__ tile_start,
                                                             unsigned long long* __restrict__ ticket,
                                                             int64_t target_units)
{
    __shared__ int64_t s_scan[33];
    __shared__ int64_t s_target;
    const int tid = threadIdx.x;
    if (SKINNY) {
        // doubles moved by all streaming blocks / target_units = doubles per unit (see sk_ucols)
        int64_t carry = 0;
        for (int64_t base = 0; base < nblocks; base += 1024) {
            const int64_t i = base + tid;
            int64_t v = 0;
            if (i < nblocks) {
                const int64_t blk = subset ? subset[i] : i;
                const int64_t M = out_mn[2 * blk], N = out_mn[2 * blk + 1];
                if (M > 0 && M <= SK_MAXM && N > 0) {
                    int64_t J = 0;
                    for (int64_t q = out_pair_ptr[blk]; q < out_pair_ptr[blk + 1]; ++q) J += pairs[3 * q + 2];
                    v = N * (J + M);
                }
            }
            int64_t total;
            block_exclusive_scan<int64_t>(v, s_scan, &total);
            carry += total;
        }
        if (tid == 0) {
            const int64_t t = (carry + target_units - 1) / (target_units > 0 ? target_units : 1);
            s_target = t > SK_CHUNK ? t : SK_CHUNK;
        }
        __syncthreads();
    }
    for (int64_t i = tid; i < nblocks; i += 1024) {
        const int64_t blk = subset ? subset[i] : i;
        double ksum = 0;
        for (int64_t q = out_pair_ptr[blk]; q < out_pair_ptr[blk + 1]; ++q) ksum += (double)pairs[3 * q + 2];
        const int64_t nt = SKINNY ? sk_units(out_mn[2 * blk], out_mn[2 * blk + 1], sk_ucols(s_target, (int64_t)ksum, out_mn[2 * blk]))
                                  : n_units<BM, BN, SKINNY>(out_mn[2 * blk], out_mn[2 * blk + 1]);
        // per-tile cost decides the order (long k-loops first), ties keep table order
        cost[i] = (nt > 0) ? ksum : -1.0;
    }
    __syncthreads();
    for (int64_t i = tid; i < nblocks; i += 1024) {
        const double ci = cost[i];
        int64_t rank = 0;
        for (int64_t j = 0; j < nblocks; ++j) {
            const double cj = cost[j];
            rank += (cj > ci) || (cj == ci && j < i);
        }
        order[rank] = subset ? subset[i] : i;
    }
    __syncthreads();
    int64_t carry = 0;
    for (int64_t base = 0; base < nblocks; base += 1024) {
        const int64_t i = base + tid;
        int64_t v = 0;
        if (i < nblocks) {
            const int64_t blk = order[i];
            if (SKINNY) {
                int64_t J = 0;
                for (int64_t q = out_pair_ptr[blk]; q < out_pair_ptr[blk + 1]; ++q) J += pairs[3 * q + 2];
                v = sk_units(out_mn[2 * blk], out_mn[2 * blk + 1], sk_ucols(s_target, J, out_mn[2 * blk]));
            } else v = n_units<BM, BN, SKINNY>(out_mn[2 * blk], out_mn[2 * blk + 1]);
        }
        int64_t total;
        int64_t pos = block_exclusive_scan<int64_t>(v, s_scan, &total);
        if (i < nblocks) tile_start[i] = carry + pos;
        carry += total;
    }
    if (tid == 0) {
        tile_start[nblocks] = carry;
        *ticket = SKINNY ? (unsigned long long)s_target : 0ull;   // the streaming kernel reads its doubles-per-unit target here
        if (SKINNY) ticket[1] = 0ull;                             // ... and draws its units from this counter
    }
}

// Warp (wm, wn) owns the contiguous WM x WN sub-tile = MI x NI groups of 8 x 8; of a ragged tile it computes only the groups
// that reach into the block (compute_stage_edge)
template <bool CPLX, int BM_, int BN_, int BK_, int WARPS_M_, int WARPS_N_, int STAGES_, int MINB_ = 1>
struct GemmCfg {
    static constexpr int MINB = MINB_;      // resident CTAs per SM (persistent grid = MINB x #SMs)
    static constexpr bool cplx = CPLX;
    static constexpr int BM = BM_, BN = BN_, BK = BK_, WARPS_M = WARPS_M_, WARPS_N = WARPS_N_, STAGES = STAGES_;
    static constexpr int THREADS = WARPS_M * WARPS_N * 32;
    static constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
    static constexpr int MI = WM / 8, NI = WN / 8;
    static constexpr int ES = CPLX ? 2 : 1;                 // doubles per element
    // row strides in ELEMENTS; bank-conflict-free fragment loads need
    //   real:    lda = 4 (mod 16) doubles,  ldb = 4 (mod 16)
    //   complex: lda = 4 (mod 8) elements,  ldb = 2 (mod 8)
    static constexpr int LDA = (BK + 15) / 16 * 16 + 4;
    static constexpr int LDB = CPLX ? BN + 2 : BN + 4;
    static constexpr int A_STAGE = BM * LDA * ES;           // doubles
    static constexpr int B_STAGE = BK * LDB * ES;           // doubles
    static constexpr size_t SMEM = (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double);
    static_assert(BK % 4 == 0 && WM % 8 == 0 && WN % 8 == 0, "tile shape");
    static_assert(CPLX ? (LDA % 8 == 4 && LDB % 8 == 2) : (LDA % 16 == 4 && LDB % 16 == 4), "padding");
};

// Contiguous warp tiles: all k-steps of one stage for a warp of which only the first MIE x NJE 8 x 8 groups lie
// inside the block.  The counts are template arguments, so the ragged variants are as free of predicates as the full one
// (uniform `if (i >= mi_cnt) continue` predicates inside the k-steps cost more than the DMMAs they saved); the warp picks
// its variant with one uniform switch per pipeline stage.  At chi = 4096 this takes the executed / algorithmic DMMA ratio
// of the 64 x 64 tiling from 1.11 to the 8-row granularity of the instruction.
template <class Cfg, int MIE, int NJE>
__device__ __forceinline__ void compute_stage_edge(const double* __restrict__ tA, const double* __restrict__ tB,
                                                   double (&acc)[Cfg::cplx ? 2 : 1][Cfg::MI][Cfg::NI][2], int wm, int wn, int lr, int lc)
{
    constexpr int BK = Cfg::BK, LDA = Cfg::LDA, LDB = Cfg::LDB, MI = Cfg::MI, NI = Cfg::NI;
    constexpr bool CPLX = Cfg::cplx;
#pragma unroll
    for (int kk = 0; kk < BK; kk += 4) {
        if (!CPLX) {
            double a[MIE], b[NJE];
#pragma unroll
            for (int i = 0; i < MIE; ++i) a[i] = tA[(size_t)((wm * MI + i) * 8 + lr) * LDA + kk + lc];
#pragma unroll
            for (int j = 0; j < NJE; ++j) b[j] = tB[(size_t)(kk + lc) * LDB + (wn * NI + j) * 8 + lr];
#pragma unroll
            for (int i = 0; i < MIE; ++i)
#pragma unroll
                for (int j = 0; j < NJE; ++j) dmma884(acc[0][i][j][0], acc[0][i][j][1], a[i], b[j]);
        } else {
            double2 a[MIE], b[NJE];
#pragma unroll
            for (int i = 0; i < MIE; ++i)
                a[i] = *reinterpret_cast<const double2*>(tA + ((size_t)((wm * MI + i) * 8 + lr) * LDA + kk + lc) * 2);
#pragma unroll
            for (int j = 0; j < NJE; ++j)
                b[j] = *reinterpret_cast<const double2*>(tB + ((size_t)(kk + lc) * LDB + (wn * NI + j) * 8 + lr) * 2);
#pragma unroll
            for (int i = 0; i < MIE; ++i) {
                const double nai = neg_bits(a[i].y);
#pragma unroll
                for (int j = 0; j < NJE; ++j) {
                    dmma884(acc[0][i][j][0], acc[0][i][j][1], a[i].x, b[j].x);
                    dmma884(acc[0][i][j][0], acc[0][i][j][1], nai, b[j].y);
                    dmma884(acc[CPLX ? 1 : 0][i][j][0], acc[CPLX ? 1 : 0][i][j][1], a[i].x, b[j].y);
                    dmma884(acc[CPLX ? 1 : 0][i][j][0], acc[CPLX ? 1 : 0][i][j][1], a[i].y, b[j].x);
                }
            }
        }
    }
}

template <class Cfg, int MIE>
__device__ __forceinline__ void compute_stage_edge_n(const double* __restrict__ tA, const double* __restrict__ tB,
                                                     double (&acc)[Cfg::cplx ? 2 : 1][Cfg::MI][Cfg::NI][2], int wm, int wn, int lr, int lc, int nj)
{
    static_assert(Cfg::MI == 4 && Cfg::NI == 4, "compute_stage_edge_n: 4 x 4 groups per warp");
    switch (nj) {
    case 0: break;
    case 1: compute_stage_edge<Cfg, MIE, 1>(tA, tB, acc, wm, wn, lr, lc); break;
    case 2: compute_stage_edge<Cfg, MIE, 2>(tA, tB, acc, wm, wn, lr, lc); break;
    case 3: compute_stage_edge<Cfg, MIE, 3>(tA, tB, acc, wm, wn, lr, lc); break;
    default: compute_stage_edge<Cfg, MIE, 4>(tA, tB, acc, wm, wn, lr, lc); break;
    }
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::MINB)
grouped_gemm_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C,
                    GemmTables T)
{
    constexpr int BM = Cfg::BM, BN = Cfg::BN, BK = Cfg::BK, STAGES = Cfg::STAGES;
    constexpr int ES = Cfg::ES, LDA = Cfg::LDA, LDB = Cfg::LDB, MI = Cfg::MI, NI = Cfg::NI;
    constexpr int THREADS = Cfg::THREADS;
    constexpr bool CPLX = Cfg::cplx;
    extern __shared__ __align__(16) double smem[];
    double* sA = smem;
    double* sB = smem + (size_t)STAGES * Cfg::A_STAGE;
    __shared__ long long s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp / Cfg::WARPS_N, wn = warp % Cfg::WARPS_N;      // contiguous WM x WN warp tiles
    const int lr = lane >> 2, lc = lane & 3;
    const int64_t total_tiles = T.tile_start[T.nblocks];

    for (;;) {
        __syncthreads();
        if (tid == 0) s_tile = (long long)atomicAdd(T.ticket, 1ull);
        __syncthreads();
        const int64_t tile = s_tile;
        if (tile >= total_tiles) break;
        // locate the output block of this tile
        int64_t lo = 0, hi = T.nblocks;
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (T.tile_start[mid] <= tile) lo = mid; else hi = mid;
        }
        const int64_t blk = T.order[lo];
        const int64_t M = T.out_mn[2 * blk], N = T.out_mn[2 * blk + 1];
        const int64_t tiles_n = (N + BN - 1) / BN;
        const int64_t local = tile - T.tile_start[lo];
        const int64_t m0 = (local / tiles_n) * BM, n0 = (local % tiles_n) * BN;
        const int64_t p_begin = T.out_pair_ptr[blk], p_end = T.out_pair_ptr[blk + 1];
        // 8 x 8 groups of this warp inside the block
        const int rem_m = (int)min((int64_t)BM, M - m0) - wm * Cfg::WM, rem_n = (int)min((int64_t)BN, N - n0) - wn * Cfg::WN;
        const int mi_eff = rem_m <= 0 ? 0 : min(MI, (rem_m + 7) >> 3), nj_eff = rem_n <= 0 ? 0 : min(NI, (rem_n + 7) >> 3);
        int64_t n_iters = 0;
        for (int64_t p = p_begin; p < p_end; ++p) n_iters += (T.pairs[3 * p + 2] + BK - 1) / BK;

        // ---- loader state (flattened over pairs x k-chunks) ----
        int64_t ld_p = p_begin, ld_k = 0;
        int64_t ld_aoff = 0, ld_boff = 0, ld_K = 0;
        if (ld_p < p_end) { ld_aoff = T.pairs[3 * ld_p]; ld_boff = T.pairs[3 * ld_p + 1]; ld_K = T.pairs[3 * ld_p + 2]; }
        const bool vecB = (N % 2 == 0) || CPLX;

        auto issue_load = [&](int stage) {
            if (ld_p < p_end) {
                double* dA = sA + (size_t)stage * Cfg::A_STAGE;
                double* dB = sB + (size_t)stage * Cfg::B_STAGE;
                const int64_t K = ld_K, k0 = ld_k;
                const double* gA = A + (ld_aoff + m0 * K + k0) * ES;   // element (m0, k0)
                const double* gB = B + (ld_boff + k0 * N + n0) * ES;   // element (k0, n0)
                // per-thread invariants (chunk column, first row) are hoisted by load_tile: ~7 instructions per copy
                // instead of a divide / modulo / 64-bit multiply per chunk (the loader used to be 80 % of the
                // instructions of a pipeline stage and kept the DMMA pipe at 73 %)
                const int rows_a = (int)min((int64_t)BM, M - m0), rows_b = (int)min((int64_t)BK, K - k0);
                const int cols_a = (int)min((int64_t)BK, K - k0) * ES, cols_b = (int)min((int64_t)BN, N - n0) * ES;
#ifndef CYB_GEMM_NOLOAD   /* experiment: ceiling of the consumer side (results are garbage) */
                if (CPLX) {
                    // one complex element = 16 bytes, always aligned
                    load_tile<BM, BK, 2, LDA * 2, THREADS>(dA, gA, K * 2, rows_a, cols_a, A, tid);
                    load_tile<BK, BN, 2, LDB * 2, THREADS>(dB, gB, N * 2, rows_b, cols_b, B, tid);
                } else {
                    if ((K & 1) == 0) load_tile<BM, BK / 2, 2, LDA, THREADS>(dA, gA, K, rows_a, cols_a, A, tid);
                    else load_tile<BM, BK, 1, LDA, THREADS>(dA, gA, K, rows_a, cols_a, A, tid);
                    if (vecB) load_tile<BK, BN / 2, 2, LDB, THREADS>(dB, gB, N, rows_b, cols_b, B, tid);
                    else load_tile<BK, BN, 1, LDB, THREADS>(dB, gB, N, rows_b, cols_b, B, tid);
                }
#endif
                ld_k += BK;
                if (ld_k >= K) {
                    ++ld_p; ld_k = 0;
                    if (ld_p < p_end) { ld_aoff = T.pairs[3 * ld_p]; ld_boff = T.pairs[3 * ld_p + 1]; ld_K = T.pairs[3 * ld_p + 2]; }
                }
            }
            cp_async_commit();
        };

        // accumulators: real [MI][NI][2]; complex re and im planes
        double acc[CPLX ? 2 : 1][MI][NI][2];
#pragma unroll
        for (int z = 0; z < (CPLX ? 2 : 1); ++z)
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NI; ++j) { acc[z][i][j][0] = 0.0; acc[z][i][j][1] = 0.0; }

#pragma unroll
        for (int s = 0; s < STAGES - 1; ++s) issue_load(s);

        int stage = 0, stage_ld = STAGES - 1;     // rolling stage indices (no 64-bit modulo in the loop)
        for (int64_t it = 0; it < n_iters; ++it, stage = stage + 1 == STAGES ? 0 : stage + 1) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            issue_load(stage_ld);
            stage_ld = stage_ld + 1 == STAGES ? 0 : stage_ld + 1;
            const double* tA = sA + (size_t)stage * Cfg::A_STAGE;
            const double* tB = sB + (size_t)stage * Cfg::B_STAGE;
            // the number of live groups is hoisted out of the unrolled k-steps: one uniform switch per stage, every variant
            // free of predicates, so the fragment loads of the next k-step are scheduled between the DMMAs of the current one
            if (mi_eff == MI && nj_eff == NI) compute_stage_edge<Cfg, MI, NI>(tA, tB, acc, wm, wn, lr, lc);     // the common case first
            else switch (mi_eff) {
            case 0: break;
            case 1: compute_stage_edge_n<Cfg, 1>(tA, tB, acc, wm, wn, lr, lc, nj_eff); break;
            case 2: compute_stage_edge_n<Cfg, 2>(tA, tB, acc, wm, wn, lr, lc, nj_eff); break;
            case 3: compute_stage_edge_n<Cfg, 3>(tA, tB, acc, wm, wn, lr, lc, nj_eff); break;
            default: compute_stage_edge_n<Cfg, 4>(tA, tB, acc, wm, wn, lr, lc, nj_eff); break;
            }
        }
        cp_async_wait<0>();

        // ---- epilogue: registers -> global ----
        double* Cb = C + T.out_off[blk] * ES;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            const int64_t r = m0 + (wm * MI + i) * 8 + lr;
            if (r >= M) continue;
#pragma unroll
            for (int j = 0; j < NI; ++j) {
                const int64_t c = n0 + (wn * NI + j) * 8 + lc * 2;
                if (!CPLX) {
                    double* p = Cb + r * N + c;
                    if (c + 1 < N && ((N & 1) == 0)) {
                        *reinterpret_cast<double2*>(p) = make_double2(acc[0][i][j][0], acc[0][i][j][1]);
                    } else {
                        if (c < N) p[0] = acc[0][i][j][0];
                        if (c + 1 < N) p[1] = acc[0][i][j][1];
                    }
                } else {
                    double* p = Cb + (r * N + c) * 2;
                    if (c < N) *reinterpret_cast<double2*>(p) = make_double2(acc[0][i][j][0], acc[CPLX ? 1 : 0][i][j][0]);
                    if (c + 1 < N) *reinterpret_cast<double2*>(p + 2) = make_double2(acc[0][i][j][1], acc[CPLX ? 1 : 0][i][j][1]);
                }
            }
        }
    }
}


// ---- streaming kernel for blocks with M <= SK_MAXM -------------------------------------------------
// C (M x N) = sum_p A_p (M x K_p) B_p (K_p x N): every thread owns VEC adjacent columns of one chunk and
// keeps the M partial sums in registers; B is read exactly once (coalesced along N), A_p is a handful of
// values broadcast through L1.  Algorithmic traffic = |B| + |C|; bound by HBM bandwidth.
template <bool CPLX, int VEC, int MM>
__device__ __noinline__ void skinny_unit_mm(const double* __restrict__ A, const double* __restrict__ B,
                                            double* __restrict__ Cb, const int64_t* __restrict__ pairs,
                                            int64_t p_begin, int64_t p_end, int M, int64_t N, int64_t n)
{
    constexpr int W = (CPLX ? 2 : 1) * VEC;   // doubles per thread per row
    double acc[MM][W];
#pragma unroll
    for (int m = 0; m < MM; ++m)
#pragma unroll
        for (int w = 0; w < W; ++w) acc[m][w] = 0.0;
    for (int64_t p = p_begin; p < p_end; ++p) {
        const int64_t a_off = pairs[3 * p], b_off = pairs[3 * p + 1];
        const int K = (int)pairs[3 * p + 2];
        const double* __restrict__ Ap = A + a_off * (CPLX ? 2 : 1);
        const double* __restrict__ Bp = B + (b_off + n) * (CPLX ? 2 : 1);
        // batches of 4 k-rows: all loads of a batch are issued before the first FMA, so every thread keeps
        // up to 4 * W * 8 bytes in flight (an in-order warp would otherwise stall on each load in turn)
        for (int k0 = 0; k0 < K; k0 += 4) {
            double b[4][W];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                if (k0 + kk < K) {
                    const double* row = Bp + (int64_t)(k0 + kk) * N * (CPLX ? 2 : 1);
                    if (W == 1) b[kk][0] = __ldcs(row);
                    else {
#pragma unroll
                        for (int w = 0; w < W; w += 2) {
                            const double2 t = __ldcs(reinterpret_cast<const double2*>(row + w));
                            b[kk][w] = t.x; b[kk][w + 1 < W ? w + 1 : w] = t.y;
                        }
                    }
                }
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                if (k0 + kk < K) {
                    const int k = k0 + kk;
#pragma unroll
                    for (int m = 0; m < MM; ++m) {
                        if (m < M) {
                            if (!CPLX) {
                                const double a = __ldg(Ap + m * K + k);
#pragma unroll
                                for (int w = 0; w < W; ++w) acc[m][w] = fma(a, b[kk][w], acc[m][w]);
                            } else {
                                const double2 a = __ldg(reinterpret_cast<const double2*>(Ap) + m * K + k);
#pragma unroll
                                for (int v = 0; v < VEC; ++v) {
                                    acc[m][2 * v] = fma(a.x, b[kk][2 * v], fma(-a.y, b[kk][2 * v + 1], acc[m][2 * v]));
                                    acc[m][2 * v + 1] = fma(a.x, b[kk][2 * v + 1], fma(a.y, b[kk][2 * v], acc[m][2 * v + 1]));
                                }
                            }
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int m = 0; m < MM; ++m) {
        if (m < M) {
            double* c = Cb + ((int64_t)m * N + n) * (CPLX ? 2 : 1);
            if (W == 1) __stcs(c, acc[m][0]);
            else {
#pragma unroll
                for (int w = 0; w < W; w += 2) __stcs(reinterpret_cast<double2*>(c + w), make_double2(acc[m][w], acc[m][w + 1]));
            }
        }
    }
}

template <bool CPLX, int VEC>
__device__ __forceinline__ void skinny_unit(const double* __restrict__ A, const double* __restrict__ B,
                                            double* __restrict__ Cb, const int64_t* __restrict__ pairs,
                                            int64_t p_begin, int64_t p_end, int M, int64_t N, int64_t n)
{
    if (M <= SK_MAXM / 2) skinny_unit_mm<CPLX, VEC, SK_MAXM / 2>(A, B, Cb, pairs, p_begin, p_end, M, N, n);
    else if (VEC == 4) {   // 8 rows x 4 columns of accumulators would spill: two passes of 2 columns
        skinny_unit_mm<CPLX, 2, SK_MAXM>(A, B, Cb, pairs, p_begin, p_end, M, N, n);
        skinny_unit_mm<CPLX, 2, SK_MAXM>(A, B, Cb, pairs, p_begin, p_end, M, N, n + 2);
    } else skinny_unit_mm<CPLX, VEC, SK_MAXM>(A, B, Cb, pairs, p_begin, p_end, M, N, n);
}

// Fast path of the streaming kernel.  All (pair, k) rows of the output block are flattened once per work
// unit into a shared-memory row table (element offset of B_p[k, 0]) and a coefficient table
// (A_p[m, k]); the streaming loop is then free of pointer chasing: per batch of RB rows every thread issues
// its RB * W / 2 16-byte loads back to back, then does the FMAs with coefficients broadcast from shared memory.
constexpr int SK_J = 64;     // rows the tables hold; blocks with more rows take the generic path

// Streaming loop of the fast path.  Every thread owns CW adjacent columns at each of the unit's column
// positions (stride PS) and a PRIVATE ring of SK_RING 16-byte slots in shared memory that it fills with
// cp.async: a step = SR table rows of one position, D = SK_RING / SR steps are in flight per thread (up to
// 256 bytes, 768 threads per SM), independent of how few rows J a block has -- with plain loads a block
// with J = 1 .. 4 rows kept only J * 16 bytes per thread in flight and the kernel sat at 30 % of HBM speed.
// Nobody else reads a thread's slots, so there is no barrier in the loop, only cp.async.wait_group.
constexpr int SK_RING = 16;
constexpr size_t SK_SMEM_RING = (size_t)SK_RING * SK_THREADS * 16;

__device__ __forceinline__ void sk_cp(void* smem, const void* g, int bytes, bool pred)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? bytes : 0;
    if (bytes == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(sa), "l"(g), "r"(sz) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(sa), "l"(g), "r"(sz) : "memory");
}

// CPLX: one complex element per slot; real: VEC2 ? two columns (16-byte slot) : one column (8-byte slot)
template <bool CPLX, bool VEC2, int MM, int SR>
__device__ __noinline__ void skinny_stream(const double* __restrict__ B, double* __restrict__ Cb,
                                           const int64_t* s_row, const double* s_coef, double2* ring,
                                           int J, int64_t N, int64_t c0, int npos)
{
    constexpr int D = SK_RING / SR;                 // steps in flight
    constexpr int CW = (!CPLX && VEC2) ? 2 : 1;     // columns per thread per position
    constexpr int SB = (CPLX || VEC2) ? 16 : 8;     // bytes per slot
    constexpr int ES = CPLX ? 2 : 1;
    constexpr int W = CPLX ? 2 : CW;                // doubles per slot
    const int tid = threadIdx.x;
    const int nb = (J + SR - 1) / SR;               // steps per position
    const int nsteps = npos * nb;
    const int64_t n0 = c0 + (int64_t)tid * CW;
    double2* my = ring + tid;                       // slot q of this thread: my[q * SK_THREADS]
    auto issue = [&](int step) {
        if (step < nsteps) {
            const int pos = step / nb, bt = step - pos * nb;
            const int64_t n = n0 + (int64_t)pos * (SK_THREADS * CW);
            const bool inb = n < N;
#pragma unroll
            for (int r = 0; r < SR; ++r) {
                const int j = bt * SR + r;
                const bool ok = inb && j < J;
                const double* src = ok ? B + (s_row[j] + n) * ES : B;
                sk_cp(my + ((step % D) * SR + r) * SK_THREADS, src, SB, ok);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
#pragma unroll 1
    for (int st = 0; st < D - 1; ++st) issue(st);
    double acc[MM][W];
#pragma unroll 1
    for (int step = 0; step < nsteps; ++step) {
        issue(step + D - 1);
        asm volatile("cp.async.wait_group %0;\n" ::"n"(D - 1) : "memory");
        const int pos = step / nb, bt = step - pos * nb;
        if (bt == 0) {
#pragma unroll
            for (int m = 0; m < MM; ++m)
#pragma unroll
                for (int w = 0; w < W; ++w) acc[m][w] = 0.0;
        }
#pragma unroll
        for (int r = 0; r < SR; ++r) {
            const int j = bt * SR + r;          // slots beyond J were zero-filled
            const int jj = j < J ? j : J - 1;
            const double2 t = my[((step % D) * SR + r) * SK_THREADS];
#pragma unroll
            for (int m = 0; m < MM; ++m) {
                if (!CPLX) {
                    const double a = s_coef[m * SK_J + jj];
                    acc[m][0] = fma(a, t.x, acc[m][0]);
                    if (W == 2) acc[m][W - 1] = fma(a, t.y, acc[m][W - 1]);
                } else {
                    const double ax = s_coef[(m * SK_J + jj) * 2], ay = s_coef[(m * SK_J + jj) * 2 + 1];
                    acc[m][0] = fma(ax, t.x, fma(-ay, t.y, acc[m][0]));
                    acc[m][W - 1] = fma(ax, t.y, fma(ay, t.x, acc[m][W - 1]));
                }
            }
        }
        if (bt == nb - 1) {
            const int64_t n = n0 + (int64_t)pos * (SK_THREADS * CW);
            if (n < N) {
#pragma unroll
                for (int m = 0; m < MM; ++m) {
                    double* c = Cb + ((int64_t)m * N + n) * ES;
                    if (W == 1) __stcs(c, acc[m][0]);
                    else __stcs(reinterpret_cast<double2*>(c), make_double2(acc[m][0], acc[m][W - 1]));
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

template <bool CPLX, bool VEC2, int MM>
__device__ __forceinline__ void skinny_stream_sr(const double* __restrict__ B, double* __restrict__ Cb,
                                                 const int64_t* s_row, const double* s_coef, double2* ring,
                                                 int J, int64_t N, int64_t c0, int npos)
{
    if (J == 1) skinny_stream<CPLX, VEC2, MM, 1>(B, Cb, s_row, s_coef, ring, J, N, c0, npos);
    else if (J == 2) skinny_stream<CPLX, VEC2, MM, 2>(B, Cb, s_row, s_coef, ring, J, N, c0, npos);
    else if (J <= 4) skinny_stream<CPLX, VEC2, MM, 4>(B, Cb, s_row, s_coef, ring, J, N, c0, npos);
    else skinny_stream<CPLX, VEC2, MM, 8>(B, Cb, s_row, s_coef, ring, J, N, c0, npos);
}

template <bool CPLX, bool VEC2>
__device__ __forceinline__ void skinny_stream_m(const double* __restrict__ B, double* __restrict__ Cb,
                                                const int64_t* s_row, const double* s_coef, double2* ring,
                                                int J, int M, int64_t N, int64_t c0, int npos)
{
    switch (M) {   // uniform over the CTA
    case 1: skinny_stream_sr<CPLX, VEC2, 1>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 2: skinny_stream_sr<CPLX, VEC2, 2>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 3: skinny_stream_sr<CPLX, VEC2, 3>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 4: skinny_stream_sr<CPLX, VEC2, 4>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 5: skinny_stream_sr<CPLX, VEC2, 5>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 6: skinny_stream_sr<CPLX, VEC2, 6>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    case 7: skinny_stream_sr<CPLX, VEC2, 7>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    default: skinny_stream_sr<CPLX, VEC2, 8>(B, Cb, s_row, s_coef, ring, J, N, c0, npos); break;
    }
}

template <bool CPLX>
__global__ void __launch_bounds__(SK_THREADS, 3) skinny_gemm_kernel(const double* __restrict__ A,
                                                                 const double* __restrict__ B,
                                                                 double* __restrict__ C, GemmTables T)
{
    // the unit -> block lookup runs on a shared-memory copy of the prefix table (a dependent chain of
    // global loads per unit would cost as much as streaming the unit)
    constexpr int SK_CACHE = 1024;
    constexpr int ES = CPLX ? 2 : 1;
    constexpr int SK_PMAX = 32;                 // pairs the fast path handles
    __shared__ int64_t s_start[SK_CACHE + 1];
    __shared__ int64_t s_row[SK_J];
    __shared__ double s_coef[SK_MAXM * SK_J * ES];
    __shared__ int64_t s_pa[SK_PMAX], s_pb[SK_PMAX];
    __shared__ int s_pk[SK_PMAX], s_pj[SK_PMAX + 1];
    __shared__ int s_J;
    __shared__ long long s_unit;
    extern __shared__ __align__(16) double2 ring[];   // SK_RING x SK_THREADS private slots
    const bool cached = T.nblocks <= SK_CACHE;
    if (cached) for (int64_t i = threadIdx.x; i <= T.nblocks; i += SK_THREADS) s_start[i] = T.tile_start[i];
    __syncthreads();
    const int64_t* start = cached ? s_start : T.tile_start;
    const int64_t total_units = start[T.nblocks];
    const int64_t target = (int64_t)T.ticket[0];   // doubles per unit chosen by the schedule kernel
    const int lane = threadIdx.x & 31;
    for (;;) {
        __syncthreads();                            // the previous unit is done with the tables
        if (threadIdx.x == 0) s_unit = (long long)atomicAdd(T.ticket + 1, 1ull);
        __syncthreads();
        const int64_t unit = s_unit;
        if (unit >= total_units) break;
        int64_t lo = 0, hi = T.nblocks;
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (start[mid] <= unit) lo = mid; else hi = mid;
        }
        const int64_t blk = T.order[lo];
        const int M = (int)T.out_mn[2 * blk];
        const int64_t N = T.out_mn[2 * blk + 1];
        const int64_t p_begin = T.out_pair_ptr[blk], p_end = T.out_pair_ptr[blk + 1];
        double* Cb = C + T.out_off[blk] * ES;
        const bool vec2 = !CPLX && (N & 1) == 0;   // real with even N: rows stay 16-byte aligned
        // ---- pair table (warp 0), then row / coefficient tables (all threads) ----
        const int np = (int)(p_end - p_begin);
        if (threadIdx.x < 32) {
            int J = SK_J + 1;
            if (np <= SK_PMAX) {
                int64_t a_off = 0, b_off = 0;
                int K = 0;
                if (lane < np) {
                    const int64_t* pr = T.pairs + 3 * (p_begin + lane);
                    a_off = pr[0]; b_off = pr[1]; K = (int)min(pr[2], (int64_t)(SK_J + 1));
                }
                int incl = K;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int o = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += o;
                }
                J = __shfl_sync(0xffffffffu, incl, 31);
                s_pa[lane] = a_off; s_pb[lane] = b_off; s_pk[lane] = K; s_pj[lane] = incl - K;
            }
            if (lane == 0) s_J = J;
        }
        __syncthreads();
        const int J = s_J;
        const bool fast = J <= SK_J;
        if (fast) {
            for (int idx = threadIdx.x; idx < J * (M + 1); idx += SK_THREADS) {
                const int j = idx % J, m = idx / J;
                int pi = 0;
                while (pi + 1 < np && s_pj[pi + 1] <= j) ++pi;
                const int k = j - s_pj[pi], K = s_pk[pi];
                if (m == M) s_row[j] = s_pb[pi] + (int64_t)k * N;
                else if (!CPLX) s_coef[m * SK_J + j] = __ldg(A + s_pa[pi] + m * K + k);
                else {
                    const double2 v = __ldg(reinterpret_cast<const double2*>(A) + s_pa[pi] + m * K + k);
                    s_coef[(m * SK_J + j) * 2] = v.x;
                    s_coef[(m * SK_J + j) * 2 + 1] = v.y;
                }
            }
        }
        __syncthreads();
        int64_t Jtot = J;
        if (!fast) { Jtot = 0; for (int64_t q = p_begin; q < p_end; ++q) Jtot += T.pairs[3 * q + 2]; }
        const int64_t ucols = sk_ucols(target, Jtot, M);
        const int npass = (int)(ucols / SK_CHUNK);
        const int64_t u0 = (unit - start[lo]) * ucols;
        if (fast) {
            // column positions of this unit that start inside the block
            const int64_t cols = (N - u0 < ucols) ? N - u0 : ucols;
            if (vec2) skinny_stream_m<CPLX, true>(B, Cb, s_row, s_coef, ring, J, M, N, u0, (int)((cols + SK_CHUNK - 1) / SK_CHUNK));
            else skinny_stream_m<CPLX, false>(B, Cb, s_row, s_coef, ring, J, M, N, u0, (int)((cols + SK_THREADS - 1) / SK_THREADS));
            continue;
        }
#pragma unroll 1
        for (int pass = 0; pass < npass; ++pass) {
            const int64_t c0 = u0 + (int64_t)pass * SK_CHUNK;
            if (c0 >= N) break;
            if (vec2) {
                const int64_t n = c0 + (int64_t)threadIdx.x * 2;
                if (n < N) skinny_unit<CPLX, CPLX ? 1 : 2>(A, B, Cb, T.pairs, p_begin, p_end, M, N, n);
            } else {
#pragma unroll 1
                for (int h = 0; h < 2; ++h) {
                    const int64_t n = c0 + threadIdx.x + h * SK_THREADS;
                    if (n < N) skinny_unit<CPLX, 1>(A, B, Cb, T.pairs, p_begin, p_end, M, N, n);
                }
            }
        }
    }
}

// float64: 64 x 64 tiles, 4 warps, four CTAs per SM (22.8 TFLOP/s on the chi = 4096 H_eff GEMMs; the 128 x 128 /
// one-CTA-per-SM layout of the first version reached 18.5, and 13 further tile / pipeline variants measured within
// -2 ... -20 % of this one -- profiles/r1b, profiles/r1c)
// (tuning aid: -DCYB_GEMM_BM= ... builds another real tiling -- 4 x 4 groups per warp: BM = 32 WARPS_M, BN = 32 WARPS_N)
#ifndef CYB_GEMM_BM
#define CYB_GEMM_BM 64
#define CYB_GEMM_BN 64
#define CYB_GEMM_BK 16
#define CYB_GEMM_STAGES 3
#define CYB_GEMM_MINB 4
#endif
#define CYB_GEMM_R CYB_GEMM_BM, CYB_GEMM_BN, CYB_GEMM_BK, CYB_GEMM_BM / 32, CYB_GEMM_BN / 32, CYB_GEMM_STAGES
#define CYB_GEMM_R_MINB CYB_GEMM_MINB
using CfgR = GemmCfg<false, CYB_GEMM_R, CYB_GEMM_R_MINB>;
// complex128: 128 x 64 tiles, 8 warps, one CTA per SM (24.8 TFLOP/s real-equivalent on the chi = 8192 Hubbard tdot, 0.70 of the peak)
using CfgC = GemmCfg<true, 128, 64, 8, 4, 2, 4>;

template <class Cfg>
static int launch_gemm(const void* a, const void* b, void* c, GemmTables T, cudaStream_t stream)
{
    static PerDeviceOnce configured;
    if (configured.first())
        CYB_CHECK_CUDA(cudaFuncSetAttribute(grouped_gemm_kernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
    grouped_gemm_kernel<Cfg><<<sm_count() * Cfg::MINB, Cfg::THREADS, Cfg::SMEM, stream>>>(
        static_cast<const double*>(a), static_cast<const double*>(b), static_cast<double*>(c), T);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // namespace cyb

using namespace cyb;

extern "C" {

// workspace per pass: cost[n] (double), order[n], tile_start[n+1], ticket  -> 3n + 4 words
int64_t cyb_grouped_gemm_workspace(int64_t nblocks)
{
    return (int64_t)(sizeof(int64_t) * 2 * (3 * (size_t)(nblocks > 0 ? nblocks : 0) + 4));
}

static void gemm_tables(GemmTables T[2], double* cost[2], const int64_t* out_mn, const int64_t* out_off,
                        const int64_t* out_pair_ptr, const int64_t* pairs, int64_t nblocks, void* workspace)
{
    const size_t n = (size_t)nblocks;
    const size_t pass_words = 3 * n + 4;
    char* ws = static_cast<char*>(workspace);
    for (int pass = 0; pass < 2; ++pass) {
        char* base = ws + sizeof(int64_t) * pass_words * pass;
        cost[pass] = reinterpret_cast<double*>(base);
        T[pass].out_mn = out_mn; T[pass].out_off = out_off; T[pass].out_pair_ptr = out_pair_ptr; T[pass].pairs = pairs;
        T[pass].order = reinterpret_cast<int64_t*>(base + sizeof(int64_t) * n);
        T[pass].tile_start = reinterpret_cast<int64_t*>(base + sizeof(int64_t) * 2 * n);
        T[pass].ticket = reinterpret_cast<unsigned long long*>(base + sizeof(int64_t) * (3 * n + 2));
        T[pass].nblocks = nblocks;
    }
}

int cyb_grouped_gemm_prepare(int dtype, const int64_t* out_mn, const int64_t* out_pair_ptr,
                             const int64_t* pairs, const int64_t* block_subset, int64_t nblocks,
                             void* workspace, int64_t workspace_bytes, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "grouped_gemm_prepare: bad dtype %d", dtype);
    CYB_REQUIRE(nblocks >= 0, "grouped_gemm_prepare: nblocks < 0");
    if (nblocks == 0) return CYB_OK;
    CYB_REQUIRE(workspace && workspace_bytes >= cyb_grouped_gemm_workspace(nblocks), "grouped_gemm_prepare: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    GemmTables T[2];
    double* cost[2];
    gemm_tables(T, cost, out_mn, nullptr, out_pair_ptr, pairs, nblocks, workspace);
    // streaming kernel: aim at SK_UNITS_PER_SLOT work units per resident CTA (3 per SM)
    const int ups = 8;
    const int64_t sk_target = (int64_t)sm_count() * 3 * ups;
    if (dtype == CYB_F64) {
        gemm_schedule_kernel<CfgR::BM, CfgR::BN, false><<<1, 1024, 0, stream>>>(out_mn, out_pair_ptr, pairs, block_subset, nblocks, cost[0], (int64_t*)T[0].order, (int64_t*)T[0].tile_start, T[0].ticket, 0);
        gemm_schedule_kernel<CfgR::BM, CfgR::BN, true><<<1, 1024, 0, stream>>>(out_mn, out_pair_ptr, pairs, block_subset, nblocks, cost[1], (int64_t*)T[1].order, (int64_t*)T[1].tile_start, T[1].ticket, sk_target);
    } else {
        gemm_schedule_kernel<CfgC::BM, CfgC::BN, false><<<1, 1024, 0, stream>>>(out_mn, out_pair_ptr, pairs, block_subset, nblocks, cost[0], (int64_t*)T[0].order, (int64_t*)T[0].tile_start, T[0].ticket, 0);
        gemm_schedule_kernel<CfgC::BM, CfgC::BN, true><<<1, 1024, 0, stream>>>(out_mn, out_pair_ptr, pairs, block_subset, nblocks, cost[1], (int64_t*)T[1].order, (int64_t*)T[1].tile_start, T[1].ticket, sk_target);
    }
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int64_t cyb_grouped_gemm_skinny_rows(void) { return SK_MAXM; }

int cyb_grouped_gemm_run(const void* a, const void* b, void* c, int dtype, const int64_t* out_mn,
                         const int64_t* out_off, const int64_t* out_pair_ptr, const int64_t* pairs,
                         int64_t nblocks, void* workspace, int kernels, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "grouped_gemm_run: bad dtype %d", dtype);
    CYB_REQUIRE(nblocks >= 0, "grouped_gemm_run: nblocks < 0");
    if (nblocks == 0) return CYB_OK;
    CYB_REQUIRE(workspace != nullptr, "grouped_gemm_run: no prepared workspace");
    cudaStream_t stream = (cudaStream_t)stream_;
    GemmTables T[2];
    double* cost[2];
    gemm_tables(T, cost, out_mn, out_off, out_pair_ptr, pairs, nblocks, workspace);
    // the dynamic tile counter of the persistent DMMA kernel starts at 0 for every run
    CYB_CHECK_CUDA(cudaMemsetAsync(T[0].ticket, 0, sizeof(unsigned long long), stream));
    CYB_CHECK_CUDA(cudaMemsetAsync(T[1].ticket + 1, 0, sizeof(unsigned long long), stream));   // unit counter of the streaming kernel
    const double* A = static_cast<const double*>(a);
    const double* B = static_cast<const double*>(b);
    double* C = static_cast<double*>(c);
    static PerDeviceOnce sk_configured;
    if (sk_configured.first()) {
        CYB_CHECK_CUDA(cudaFuncSetAttribute(skinny_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SK_SMEM_RING));
        CYB_CHECK_CUDA(cudaFuncSetAttribute(skinny_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SK_SMEM_RING));
    }
    const int sk_grid = sm_count() * 3;   // resident CTAs; units are drawn from an atomic ticket
    // `kernels`: bit 0 = some block has 1 <= M <= SK_MAXM rows (streaming kernel), bit 1 = some block has more (DMMA kernel);
    // the caller knows from out_mn, and a kernel without work is not launched (13 us each under ncu)
    const bool run_sk = (kernels & 1) != 0, run_mma = (kernels & 2) != 0;
    if (dtype == CYB_F64) {
        if (run_sk) {
            skinny_gemm_kernel<false><<<sk_grid, SK_THREADS, SK_SMEM_RING, stream>>>(A, B, C, T[1]);
            CYB_LAUNCH_CHECK();
        }
        if (!run_mma) return CYB_OK;
        return launch_gemm<CfgR>(a, b, c, T[0], stream);
    }
    if (run_sk) {
        skinny_gemm_kernel<true><<<sk_grid, SK_THREADS, SK_SMEM_RING, stream>>>(A, B, C, T[1]);
        CYB_LAUNCH_CHECK();
    }
    if (!run_mma) return CYB_OK;
    return launch_gemm<CfgC>(a, b, c, T[0], stream);
}

int cyb_grouped_gemm(const void* a, const void* b, void* c, int dtype, const int64_t* out_mn,
                     const int64_t* out_off, const int64_t* out_pair_ptr, const int64_t* pairs,
                     const int64_t* block_subset, int64_t nblocks, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "grouped_gemm: bad dtype %d", dtype);
    CYB_REQUIRE(nblocks >= 0, "grouped_gemm: nblocks < 0");
    if (nblocks == 0) return CYB_OK;
    const int64_t bytes = cyb_grouped_gemm_workspace(nblocks);
    Scratch ws((size_t)bytes, (cudaStream_t)stream_);
    if (!ws.ptr) return CYB_ERR_CUDA;
    int rc = cyb_grouped_gemm_prepare(dtype, out_mn, out_pair_ptr, pairs, block_subset, nblocks, ws.ptr, bytes, stream_);
    if (rc != CYB_OK) return rc;
    return cyb_grouped_gemm_run(a, b, c, dtype, out_mn, out_off, out_pair_ptr, pairs, nblocks, ws.ptr, 3, stream_);
}

}  // extern "C"
