// (a1) charge-sector bookkeeping on device: lexsort / take / row differences / pack / merge-join,
// abelian charge fusion, and the compose planner (join + sort + match of block_inds).
// Every table is int64 row-major, exactly the reference's BlockInds layout
// (include/cyten/backends/block_inds.h:22-160); results are bit-exact with
// src/backends/block_inds.cpp and the planning part of src/backends/abelian.cpp:1238-1469.
//
// These tables are tiny (10..10^4 rows): the kernels are latency bound, so each one is a single
// launch; the O(n^2) rank sort is stable by construction and needs no scratch.
#include "common.cuh"

namespace cyb {

constexpr int BK_THREADS = 1024;

// ---- lexsort ------------------------------------------------------------------------------
// cmp of rows i, j with the LAST column most significant (np.lexsort(rows.T)).
__device__ __forceinline__ int cmp_rows_lex(const int64_t* __restrict__ ri,
                                            const int64_t* __restrict__ rj, int ncols)
{
    for (int c = ncols - 1; c >= 0; --c) {
        int64_t a = ri[c], b = rj[c];
        if (a < b) return -1;
        if (a > b) return 1;
    }
    return 0;
}

// rank sort: perm[rank(i)] = i with rank(i) = #{j : row_j < row_i or (row_j == row_i and j < i)}.
// Rows of the compared tile are staged in shared memory.
template <int TILE>
__global__ void __launch_bounds__(256) lexsort_rank_kernel(const int64_t* __restrict__ rows,
                                                           int64_t nrows, int ncols,
                                                           int64_t* __restrict__ perm)
{
    extern __shared__ int64_t s_rows[];  // TILE * ncols
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t mine[CYB_MAX_RANK * 2];
    const bool active = i < nrows;
    if (active)
        for (int c = 0; c < ncols; ++c) mine[c] = rows[i * ncols + c];
    int64_t rank = 0;
    for (int64_t base = 0; base < nrows; base += TILE) {
        const int64_t cnt = min((int64_t)TILE, nrows - base);
        __syncthreads();
        for (int64_t t = threadIdx.x; t < cnt * ncols; t += blockDim.x) s_rows[t] = rows[base * ncols + t];
        __syncthreads();
        if (active) {
            for (int64_t t = 0; t < cnt; ++t) {
                const int64_t j = base + t;
                int c = cmp_rows_lex(&s_rows[t * ncols], mine, ncols);
                rank += (c < 0) || (c == 0 && j < i);
            }
        }
    }
    if (active) perm[rank] = i;
}

__global__ void take_rows_kernel(const int64_t* __restrict__ rows, int64_t ncols,
                                 const int64_t* __restrict__ idx, int64_t nrows_out,
                                 const int64_t* __restrict__ cols, int64_t ncols_out,
                                 int64_t* __restrict__ out)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows_out * ncols_out) return;
    const int64_t r = t / ncols_out, c = t % ncols_out;
    const int64_t sr = idx ? idx[r] : r;
    const int64_t sc = cols ? cols[c] : c;
    out[t] = rows[sr * ncols + sc];
}

// single CTA; starts_out gets the start of every run, then nrows; count_out = runs + 1
__global__ void __launch_bounds__(BK_THREADS) row_differences_kernel(
    const int64_t* __restrict__ rows, int64_t nrows, int ncols, int64_t* __restrict__ starts_out,
    int64_t* __restrict__ count_out)
{
    __shared__ int64_t s_scan[33];
    int64_t carry = 0;
    for (int64_t base = 0; base < nrows; base += BK_THREADS) {
        const int64_t i = base + threadIdx.x;
        int64_t flag = 0;
        if (i < nrows) {
            if (i == 0) flag = 1;
            else {
                for (int c = 0; c < ncols; ++c)
                    if (rows[(i - 1) * ncols + c] != rows[i * ncols + c]) { flag = 1; break; }
            }
        }
        int64_t total;
        int64_t pos = block_exclusive_scan<int64_t>(flag, s_scan, &total);
        if (flag) starts_out[carry + pos] = i;
        carry += total;
    }
    if (threadIdx.x == 0) {
        starts_out[carry] = nrows;
        count_out[0] = carry + 1;
    }
}

__global__ void pack_rows_kernel(const int64_t* __restrict__ rows, int64_t nrows, int ncols,
                                 const int64_t* __restrict__ strides, int64_t* __restrict__ keys)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    int64_t s = 0;
    for (int c = 0; c < ncols; ++c) s += rows[r * ncols + c] * strides[c];
    keys[r] = s;
}

// single CTA merge-join: each a[i] binary-searches b (both strictly ascending)
__global__ void __launch_bounds__(BK_THREADS) common_sorted_kernel(
    const int64_t* __restrict__ a, int64_t na, const int64_t* __restrict__ b, int64_t nb,
    int64_t* __restrict__ pairs, int64_t* __restrict__ count_out)
{
    __shared__ int64_t s_scan[33];
    int64_t carry = 0;
    for (int64_t base = 0; base < na; base += BK_THREADS) {
        const int64_t i = base + threadIdx.x;
        int64_t flag = 0, jfound = -1;
        if (i < na) {
            const int64_t key = a[i];
            int64_t lo = 0, hi = nb;
            while (lo < hi) {
                int64_t mid = (lo + hi) >> 1;
                if (b[mid] < key) lo = mid + 1; else hi = mid;
            }
            if (lo < nb && b[lo] == key) { flag = 1; jfound = lo; }
        }
        int64_t total;
        int64_t pos = block_exclusive_scan<int64_t>(flag, s_scan, &total);
        if (flag) {
            pairs[2 * (carry + pos)] = i;
            pairs[2 * (carry + pos) + 1] = jfound;
        }
        carry += total;
    }
    if (threadIdx.x == 0) count_out[0] = carry;
}

// abelian fusion: component-wise sum, reduced mod N for Z_N factors (0 = U(1))
__global__ void fuse_charges_kernel(const int64_t* __restrict__ bi, int64_t nrows, int ncols,
                                    const int64_t* __restrict__ cols, int nlegs,
                                    const int64_t* __restrict__ sectors,
                                    const int64_t* __restrict__ sector_ptr,
                                    const int64_t* __restrict__ modulus, int nsym,
                                    int64_t* __restrict__ out)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows * nsym) return;
    const int64_t r = t / nsym;
    const int s = (int)(t % nsym);
    int64_t q = 0;
    for (int f = 0; f < nlegs; ++f) {
        const int64_t b = bi[r * ncols + cols[f]];
        q += sectors[(sector_ptr[f] + b) * nsym + s];
    }
    const int64_t m = modulus[s];
    if (m > 0) { q %= m; if (q < 0) q += m; }
    out[t] = q;
}

// ---- compose planner ----------------------------------------------------------------------
struct ComposePlanArgs {
    const int64_t* a_bi; int64_t na; int ra; int a_keep;
    const int64_t* b_bi; int64_t nb; int rb; int ncontr;
    const int64_t* contr_nsec;
    const int64_t* a_charge; const int64_t* b_charge; int nsym;
    const int64_t* a_mk; const int64_t* b_kn; const int64_t* a_off; const int64_t* b_off;
    int64_t align_elems; int64_t cap_out; int64_t cap_pairs;
    int64_t* header; int64_t* out_bi; int64_t* out_rowcol; int64_t* out_mn; int64_t* out_off;
    int64_t* out_pair_ptr; int64_t* pairs;
    // workspace (int64 each)
    int64_t* a_key;     // [na] packed contracted key (original order)
    int64_t* a_sort;    // [na] sorted position -> original row
    int64_t* a_skey;    // [na] keys in sorted order
    int64_t* b_key;     // [nb]
    int64_t* a_gstart;  // [na + 1]
    int64_t* b_gstart;  // [nb + 1]
    int64_t* gb_out;    // [nb + 1] exclusive scan of outputs per B group
    int64_t* gb_pairs;  // [nb + 1]
};

__device__ __forceinline__ int cmp_keep_a(const int64_t* __restrict__ bi, int ra, int a_keep,
                                          int64_t i, int64_t j)
{
    for (int c = a_keep - 1; c >= 0; --c) {
        int64_t x = bi[i * ra + c], y = bi[j * ra + c];
        if (x < y) return -1;
        if (x > y) return 1;
    }
    return 0;
}

// number of common entries of two ascending key lists; optionally emits them
template <bool EMIT>
__device__ __forceinline__ int64_t merge_common(const int64_t* __restrict__ ka, int64_t na,
                                                const int64_t* __restrict__ kb, int64_t nb,
                                                int64_t a_base, int64_t b_base,
                                                const ComposePlanArgs& P, int64_t pair_pos)
{
    int64_t i = 0, j = 0, n = 0;
    while (i < na && j < nb) {
        const int64_t x = ka[i], y = kb[j];
        if (x < y) ++i;
        else if (x > y) ++j;
        else {
            if (EMIT) {
                const int64_t arow = P.a_sort[a_base + i], brow = b_base + j;
                int64_t* p = P.pairs + 3 * (pair_pos + n);
                p[0] = P.a_off[arow];
                p[1] = P.b_off[brow];
                p[2] = P.a_mk[2 * arow + 1];
            }
            ++n; ++i; ++j;
        }
    }
    return n;
}

__global__ void __launch_bounds__(BK_THREADS) compose_plan_kernel(ComposePlanArgs P)
{
    __shared__ int64_t s_scan[33];
    __shared__ int64_t s_strides[CYB_MAX_RANK * 2];
    __shared__ int64_t s_nGa, s_nGb;
    const int tid = threadIdx.x;
    const int64_t na = P.na, nb = P.nb;
    const int ra = P.ra, rb = P.rb, a_keep = P.a_keep, ncontr = P.ncontr;
    const int b_keep = rb - ncontr;

    // F-style strides over the contracted legs in B's column order (abelian.cpp:1270-1283)
    if (tid == 0) {
        int64_t s = 1;
        for (int c = 0; c < ncontr; ++c) { s_strides[c] = s; s *= P.contr_nsec[c]; }
    }
    __syncthreads();
    for (int64_t i = tid; i < na; i += BK_THREADS) {
        int64_t k = 0;
        for (int c = 0; c < ncontr; ++c) k += P.a_bi[i * ra + a_keep + c] * s_strides[ncontr - 1 - c];
        P.a_key[i] = k;
    }
    for (int64_t j = tid; j < nb; j += BK_THREADS) {
        int64_t k = 0;
        for (int c = 0; c < ncontr; ++c) k += P.b_bi[j * rb + c] * s_strides[c];
        P.b_key[j] = k;
    }
    __syncthreads();

    // stable sort of A's rows by (keep columns lexsorted, then contracted key): abelian.cpp:1286-1301
    for (int64_t i = tid; i < na; i += BK_THREADS) {
        const int64_t ki = P.a_key[i];
        int64_t rank = 0;
        for (int64_t j = 0; j < na; ++j) {
            int c = cmp_keep_a(P.a_bi, ra, a_keep, j, i);
            if (c == 0) { const int64_t kj = P.a_key[j]; c = kj < ki ? -1 : (kj > ki ? 1 : 0); }
            rank += (c < 0) || (c == 0 && j < i);
        }
        P.a_sort[rank] = i;
        P.a_skey[rank] = ki;
    }
    __syncthreads();

    // group boundaries (find_row_differences on the keep columns): abelian.cpp:1303-1335
    {
        int64_t carry = 0;
        for (int64_t base = 0; base < na; base += BK_THREADS) {
            const int64_t r = base + tid;
            int64_t flag = 0;
            if (r < na)
                flag = (r == 0) ? 1 : (cmp_keep_a(P.a_bi, ra, a_keep, P.a_sort[r - 1], P.a_sort[r]) != 0);
            int64_t total;
            int64_t pos = block_exclusive_scan<int64_t>(flag, s_scan, &total);
            if (flag) P.a_gstart[carry + pos] = r;
            carry += total;
        }
        if (tid == 0) { P.a_gstart[carry] = na; s_nGa = carry; }
        carry = 0;
        for (int64_t base = 0; base < nb; base += BK_THREADS) {
            const int64_t r = base + tid;
            int64_t flag = 0;
            if (r < nb) {
                if (r == 0) flag = 1;
                else
                    for (int c = ncontr; c < rb; ++c)
                        if (P.b_bi[(r - 1) * rb + c] != P.b_bi[r * rb + c]) { flag = 1; break; }
            }
            int64_t total;
            int64_t pos = block_exclusive_scan<int64_t>(flag, s_scan, &total);
            if (flag) P.b_gstart[carry + pos] = r;
            carry += total;
        }
        if (tid == 0) { P.b_gstart[carry] = nb; s_nGb = carry; }
    }
    __syncthreads();
    const int64_t nGa = s_nGa, nGb = s_nGb;
    const int nsym = P.nsym;

    // pass 1: per B column group, count matching A row groups and block pairs
    int64_t out_carry = 0, pair_carry = 0;
    for (int64_t base = 0; base < nGb; base += BK_THREADS) {
        const int64_t gb = base + tid;
        int64_t n_out = 0, n_pairs = 0;
        if (gb < nGb) {
            const int64_t b0 = P.b_gstart[gb], b1 = P.b_gstart[gb + 1];
            const int64_t* qb = P.b_charge + b0 * nsym;
            for (int64_t ga = 0; ga < nGa; ++ga) {
                const int64_t a0 = P.a_gstart[ga], a1 = P.a_gstart[ga + 1];
                const int64_t* qa = P.a_charge + P.a_sort[a0] * nsym;
                bool eq = true;
                for (int s = 0; s < nsym; ++s) eq = eq && (qa[s] == qb[s]);
                if (!eq) continue;
                int64_t c = merge_common<false>(P.a_skey + a0, a1 - a0, P.b_key + b0, b1 - b0, a0, b0, P, 0);
                if (c > 0) { ++n_out; n_pairs += c; }
            }
        }
        int64_t tot_o, tot_p;
        int64_t po = block_exclusive_scan<int64_t>(n_out, s_scan, &tot_o);
        int64_t pp = block_exclusive_scan<int64_t>(n_pairs, s_scan, &tot_p);
        if (gb < nGb) { P.gb_out[gb] = out_carry + po; P.gb_pairs[gb] = pair_carry + pp; }
        out_carry += tot_o;
        pair_carry += tot_p;
    }
    const int64_t n_out_total = out_carry, n_pairs_total = pair_carry;
    const bool fits = n_out_total <= P.cap_out && n_pairs_total <= P.cap_pairs;
    __syncthreads();

    // pass 2: emit rows (row_a fastest inside col_b => result lexsorted, abelian.cpp:1424-1459)
    if (fits) {
        for (int64_t gb = tid; gb < nGb; gb += BK_THREADS) {
            const int64_t b0 = P.b_gstart[gb], b1 = P.b_gstart[gb + 1];
            const int64_t* qb = P.b_charge + b0 * nsym;
            int64_t o = P.gb_out[gb], p = P.gb_pairs[gb];
            for (int64_t ga = 0; ga < nGa; ++ga) {
                const int64_t a0 = P.a_gstart[ga], a1 = P.a_gstart[ga + 1];
                const int64_t arow = P.a_sort[a0];
                const int64_t* qa = P.a_charge + arow * nsym;
                bool eq = true;
                for (int s = 0; s < nsym; ++s) eq = eq && (qa[s] == qb[s]);
                if (!eq) continue;
                int64_t c = merge_common<true>(P.a_skey + a0, a1 - a0, P.b_key + b0, b1 - b0, a0, b0, P, p);
                if (c == 0) continue;
                int64_t* row = P.out_bi + o * (a_keep + b_keep);
                for (int k = 0; k < a_keep; ++k) row[k] = P.a_bi[arow * ra + k];
                for (int k = 0; k < b_keep; ++k) row[a_keep + k] = P.b_bi[b0 * rb + ncontr + k];
                P.out_rowcol[2 * o] = arow;
                P.out_rowcol[2 * o + 1] = b0;
                P.out_mn[2 * o] = P.a_mk[2 * arow];
                P.out_mn[2 * o + 1] = P.b_kn[2 * b0 + 1];
                P.out_pair_ptr[o] = p;
                ++o;
                p += c;
            }
        }
        if (tid == 0) P.out_pair_ptr[n_out_total] = n_pairs_total;
    }
    __syncthreads();

    // output arena offsets (aligned) and the algorithmic M*K*N count
    int64_t elem_carry = 0;
    double mkn = 0.0;
    if (fits) {
        for (int64_t base = 0; base < n_out_total; base += BK_THREADS) {
            const int64_t o = base + tid;
            int64_t sz = 0;
            if (o < n_out_total) {
                const int64_t M = P.out_mn[2 * o], N = P.out_mn[2 * o + 1];
                sz = (M * N + P.align_elems - 1) / P.align_elems * P.align_elems;
                for (int64_t q = P.out_pair_ptr[o]; q < P.out_pair_ptr[o + 1]; ++q)
                    mkn += (double)M * (double)N * (double)P.pairs[3 * q + 2];
            }
            int64_t total;
            int64_t pos = block_exclusive_scan<int64_t>(sz, s_scan, &total);
            if (o < n_out_total) P.out_off[o] = elem_carry + pos;
            elem_carry += total;
        }
    }
    // reduce mkn over the CTA
    __shared__ double s_red[32];
    mkn = warp_sum(mkn);
    __syncthreads();
    if ((tid & 31) == 0) s_red[tid >> 5] = mkn;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < BK_THREADS / 32; ++w) t += s_red[w];
        P.header[0] = n_out_total;
        P.header[1] = n_pairs_total;
        P.header[2] = elem_carry;
        P.header[3] = fits ? 1 : 0;
        P.header[4] = (int64_t)t;
        P.header[5] = nGa;
        P.header[6] = nGb;
        P.header[7] = 0;
    }
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int cyb_lexsort_rows(const int64_t* rows, int64_t nrows, int64_t ncols, int64_t* perm_out,
                     cyb_stream_t stream)
{
    CYB_REQUIRE(nrows >= 0 && ncols >= 0 && ncols <= 2 * CYB_MAX_RANK, "lexsort_rows: bad shape (%lld x %lld)",
                (long long)nrows, (long long)ncols);
    if (nrows == 0) return CYB_OK;
    constexpr int TILE = 256;
    const int threads = 256;
    const int64_t grid = (nrows + threads - 1) / threads;
    size_t smem = (size_t)TILE * (size_t)(ncols > 0 ? ncols : 1) * sizeof(int64_t);
    lexsort_rank_kernel<TILE><<<(unsigned)grid, threads, smem, (cudaStream_t)stream>>>(rows, nrows, (int)ncols, perm_out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_take_rows(const int64_t* rows, int64_t ncols, const int64_t* idx, int64_t nrows_out,
                  const int64_t* cols, int64_t ncols_out, int64_t* out, cyb_stream_t stream)
{
    CYB_REQUIRE(nrows_out >= 0 && ncols_out >= 0, "take_rows: bad shape");
    const int64_t n = nrows_out * ncols_out;
    if (n == 0) return CYB_OK;
    take_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rows, ncols, idx, nrows_out, cols, ncols_out, out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_row_differences(const int64_t* rows, int64_t nrows, int64_t ncols, int64_t* starts_out,
                        int64_t* count_out, cyb_stream_t stream)
{
    CYB_REQUIRE(nrows >= 0 && ncols >= 0, "row_differences: bad shape");
    row_differences_kernel<<<1, BK_THREADS, 0, (cudaStream_t)stream>>>(rows, nrows, (int)ncols, starts_out, count_out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_pack_rows(const int64_t* rows, int64_t nrows, int64_t ncols, const int64_t* strides,
                  int64_t* keys_out, cyb_stream_t stream)
{
    CYB_REQUIRE(nrows >= 0 && ncols >= 0, "pack_rows: bad shape");
    if (nrows == 0) return CYB_OK;
    pack_rows_kernel<<<(unsigned)((nrows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rows, nrows, (int)ncols, strides, keys_out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_common_sorted(const int64_t* a, int64_t na, const int64_t* b, int64_t nb,
                      int64_t* pairs_out, int64_t* count_out, cyb_stream_t stream)
{
    CYB_REQUIRE(na >= 0 && nb >= 0, "common_sorted: bad shape");
    common_sorted_kernel<<<1, BK_THREADS, 0, (cudaStream_t)stream>>>(a, na, b, nb, pairs_out, count_out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_fuse_charges(const int64_t* bi, int64_t nrows, int64_t ncols, const int64_t* cols,
                     int64_t nlegs, const int64_t* sectors, const int64_t* sector_ptr,
                     const int64_t* modulus, int64_t nsym, int64_t* charge_out,
                     cyb_stream_t stream)
{
    CYB_REQUIRE(nrows >= 0 && nsym >= 1, "fuse_charges: bad shape");
    const int64_t n = nrows * nsym;
    if (n == 0) return CYB_OK;
    fuse_charges_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        bi, nrows, (int)ncols, cols, (int)nlegs, sectors, sector_ptr, modulus, (int)nsym, charge_out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int64_t cyb_compose_plan_workspace(int64_t na, int64_t nb)
{
    // a_key, a_sort, a_skey, a_gstart(+1) ; b_key, b_gstart(+1), gb_out(+1), gb_pairs(+1)
    return (int64_t)sizeof(int64_t) * (4 * na + 4 * nb + 8);
}

int cyb_compose_plan(const int64_t* a_bi, int64_t na, int64_t ra, int64_t a_keep,
                     const int64_t* b_bi, int64_t nb, int64_t rb, int64_t ncontr,
                     const int64_t* contr_nsec, const int64_t* a_charge, const int64_t* b_charge,
                     int64_t nsym, const int64_t* a_mk, const int64_t* b_kn, const int64_t* a_off,
                     const int64_t* b_off, int64_t align_elems, int64_t cap_out, int64_t cap_pairs,
                     int64_t* header, int64_t* out_bi, int64_t* out_rowcol, int64_t* out_mn,
                     int64_t* out_off, int64_t* out_pair_ptr, int64_t* pairs, void* workspace,
                     int64_t workspace_bytes, cyb_stream_t stream)
{
    CYB_REQUIRE(na >= 1 && nb >= 1, "compose_plan: empty operand (na=%lld nb=%lld)", (long long)na, (long long)nb);
    CYB_REQUIRE(ra - a_keep == ncontr && ncontr >= 1 && ncontr <= rb && ncontr <= CYB_MAX_RANK * 2,
                "compose_plan: inconsistent leg counts");
    CYB_REQUIRE(align_elems >= 1, "compose_plan: align_elems must be >= 1");
    CYB_REQUIRE(workspace_bytes >= cyb_compose_plan_workspace(na, nb), "compose_plan: workspace too small");
    ComposePlanArgs P;
    P.a_bi = a_bi; P.na = na; P.ra = (int)ra; P.a_keep = (int)a_keep;
    P.b_bi = b_bi; P.nb = nb; P.rb = (int)rb; P.ncontr = (int)ncontr;
    P.contr_nsec = contr_nsec; P.a_charge = a_charge; P.b_charge = b_charge; P.nsym = (int)nsym;
    P.a_mk = a_mk; P.b_kn = b_kn; P.a_off = a_off; P.b_off = b_off;
    P.align_elems = align_elems; P.cap_out = cap_out; P.cap_pairs = cap_pairs;
    P.header = header; P.out_bi = out_bi; P.out_rowcol = out_rowcol; P.out_mn = out_mn;
    P.out_off = out_off; P.out_pair_ptr = out_pair_ptr; P.pairs = pairs;
    int64_t* w = static_cast<int64_t*>(workspace);
    P.a_key = w; w += na;
    P.a_sort = w; w += na;
    P.a_skey = w; w += na;
    P.a_gstart = w; w += na + 1;
    P.b_key = w; w += nb;
    P.b_gstart = w; w += nb + 1;
    P.gb_out = w; w += nb + 1;
    P.gb_pairs = w; w += nb + 1;
    compose_plan_kernel<<<1, BK_THREADS, 0, (cudaStream_t)stream>>>(P);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // extern "C"
