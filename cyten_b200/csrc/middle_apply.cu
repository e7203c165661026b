// (f3) One streaming pass for a small operator acting on the MIDDLE legs of a large tensor:
//
//     out[o][a][b] = sum_i C_g[o][i] * in[i][a][b]            for every group g, a < na_g, b < nb_g
//
// "in row" i of group g is the vector  in + in_off[i] + a * in_stride[i] + [0, nb)  (contiguous in b), likewise the
// output rows.  This is the two MPO applications of the DMRG effective Hamiltonian (LP . theta -> W1 -> W2 -> . RP,
// toycodes/tenpy_toycodes/d_dmrg.py:70-88; PlanarHEffective, :94-116) executed as ONE kernel between the two fat GEMMs:
// the first GEMM leaves x[vL, (wL p0 p1), vR], the second wants x'[vL, (p0 p1 wR), vR]; for a fixed pair of (vL, vR)
// sectors the blocks of x / x' are rows of a small dense operator (the contraction W1 . W2 over wC restricted to the
// charges that conserve the total), the same for every (a, b).  The reference -- and the unfused path here -- does
// compose + permute_legs twice, i.e. five read + write passes over the 5 x theta-sized intermediates; this kernel reads
// them once and writes them once.  HBM-bound: algorithmic bytes = (nin + nout) * na * nb * sizeof(element) per group.
#include "common.cuh"

namespace cyb {

constexpr int MID_THREADS = 128;
constexpr int MID_COLS = 2 * MID_THREADS;      // columns per work unit (two per thread)
constexpr int MID_MAX = 16;                    // rows per side of a group's operator

struct MidGroup {
    int64_t na, nb, nin, nout, desc0, coef0;
};

// MAXR: compile-time bound on the rows per side of every group's operator (8 for the D = 5, d = 2 effective Hamiltonian:
// half the registers of the general 16-row version, twice the resident CTAs and bytes in flight)
template <bool CPLX, int MAXR>
__global__ void __launch_bounds__(MID_THREADS, MAXR <= 8 ? 8 : 4)
middle_apply_kernel(const double* __restrict__ in, double* __restrict__ out, const MidGroup* __restrict__ groups,
                    const int64_t* __restrict__ desc, const double* __restrict__ coef, const int64_t* __restrict__ work0,
                    int64_t ngroups, int64_t nwork)
{
    constexpr int ES = CPLX ? 2 : 1;
    __shared__ double s_coef[MAXR * MAXR * ES];
    __shared__ int64_t s_desc[2 * MAXR][2];
    __shared__ MidGroup s_g;
    const int tid = threadIdx.x;
    int64_t cur = -1;
    // contiguous ranges of units per CTA: consecutive units share their group, whose tables are staged once
    const int64_t per = (nwork + gridDim.x - 1) / gridDim.x;
    const int64_t u0 = (int64_t)blockIdx.x * per, u1 = u0 + per < nwork ? u0 + per : nwork;
    // (a, chunk) of the current unit are carried from unit to unit: the 64-bit divide / modulo per unit and thread that
    // used to find them, with the index arithmetic around it, made this HBM-bound kernel issue-bound (ncu: 60 % issue
    // slots active at 3.2 TB/s)
    int64_t cur_end = 0, a = 0;
    int nin = 0, nout = 0, nb = 0, chunks = 1, chunk = 0;
    for (int64_t u = u0; u < u1; ++u) {
        if (cur < 0 || u >= cur_end) {
            int64_t lo = cur >= 0 ? cur : 0, hi = ngroups;
            while (hi - lo > 1) {
                const int64_t mid = (lo + hi) >> 1;
                if (work0[mid] <= u) lo = mid; else hi = mid;
            }
            __syncthreads();                     // everybody is done with the previous group's tables
            if (tid == 0) s_g = groups[lo];
            __syncthreads();
            nin = (int)s_g.nin; nout = (int)s_g.nout; nb = (int)s_g.nb;
            for (int e = tid; e < nin * nout * ES; e += MID_THREADS) s_coef[e] = coef[s_g.coef0 * ES + e];
            for (int e = tid; e < 2 * (nin + nout); e += MID_THREADS) s_desc[e >> 1][e & 1] = desc[2 * s_g.desc0 + e];
            __syncthreads();
            cur = lo;
            cur_end = work0[cur + 1];
            chunks = (nb + MID_COLS - 1) / MID_COLS;
            const int64_t local = u - work0[cur];
            a = local / chunks;
            chunk = (int)(local - a * chunks);
        }
        const int b0 = chunk * MID_COLS + tid, b1 = b0 + MID_THREADS;
        const bool ok0 = b0 < nb, ok1 = b1 < nb;
        if (!CPLX) {
            double x0[MAXR], x1[MAXR];
#pragma unroll
            for (int i = 0; i < MAXR; ++i) {
                x0[i] = x1[i] = 0.0;
                if (i < nin) {
                    const double* p = in + s_desc[i][0] + a * s_desc[i][1];
                    if (ok0) x0[i] = __ldcs(p + b0);
                    if (ok1) x1[i] = __ldcs(p + b1);
                }
            }
            for (int o = 0; o < nout; ++o) {
                double y0 = 0.0, y1 = 0.0;
                const double* c = s_coef + o * nin;
#pragma unroll
                for (int i = 0; i < MAXR; ++i)
                    if (i < nin) { y0 += c[i] * x0[i]; y1 += c[i] * x1[i]; }
                double* q = out + s_desc[nin + o][0] + a * s_desc[nin + o][1];
                if (ok0) __stcs(q + b0, y0);
                if (ok1) __stcs(q + b1, y1);
            }
        } else {
            double2 x0[MAXR], x1[MAXR];
#pragma unroll
            for (int i = 0; i < MAXR; ++i) {
                x0[i] = x1[i] = make_double2(0.0, 0.0);
                if (i < nin) {
                    const double2* p = reinterpret_cast<const double2*>(in) + s_desc[i][0] + a * s_desc[i][1];
                    if (ok0) x0[i] = __ldcs(p + b0);
                    if (ok1) x1[i] = __ldcs(p + b1);
                }
            }
            for (int o = 0; o < nout; ++o) {
                double2 y0 = make_double2(0.0, 0.0), y1 = y0;
                const double* c = s_coef + 2 * o * nin;
#pragma unroll
                for (int i = 0; i < MAXR; ++i)
                    if (i < nin) {
                        const double cr = c[2 * i], ci = c[2 * i + 1];
                        y0.x += cr * x0[i].x - ci * x0[i].y; y0.y += cr * x0[i].y + ci * x0[i].x;
                        y1.x += cr * x1[i].x - ci * x1[i].y; y1.y += cr * x1[i].y + ci * x1[i].x;
                    }
                double2* q = reinterpret_cast<double2*>(out) + s_desc[nin + o][0] + a * s_desc[nin + o][1];
                if (ok0) __stcs(q + b0, y0);
                if (ok1) __stcs(q + b1, y1);
            }
        }
        if (++chunk == chunks) { chunk = 0; ++a; }
    }
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int64_t cyb_middle_apply_max_rows(void) { return MID_MAX; }
int64_t cyb_middle_apply_unit_cols(void) { return MID_COLS; }

int cyb_middle_apply(const void* in, void* out, const int64_t* groups, const int64_t* desc, const void* coef,
                     const int64_t* work0, int64_t ngroups, int64_t nwork, int64_t max_rows, int dtype, cyb_stream_t stream_)
{
    CYB_REQUIRE(max_rows >= 0 && max_rows <= MID_MAX, "middle_apply: max_rows %lld outside [0, %d]", (long long)max_rows, MID_MAX);
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "middle_apply: bad dtype %d", dtype);
    CYB_REQUIRE(ngroups >= 0 && nwork >= 0, "middle_apply: negative counts");
    if (ngroups == 0 || nwork == 0) return CYB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool small = max_rows <= 8;
    int64_t grid = (int64_t)sm_count() * (small ? 8 : 4);
    if (grid > nwork) grid = nwork;
    const MidGroup* g = reinterpret_cast<const MidGroup*>(groups);
    const double *pi = static_cast<const double*>(in), *pc = static_cast<const double*>(coef);
    double* po = static_cast<double*>(out);
#define CYB_MID_LAUNCH(CP, MR) middle_apply_kernel<CP, MR><<<(unsigned)grid, MID_THREADS, 0, stream>>>(pi, po, g, desc, pc, work0, ngroups, nwork)
    if (dtype == CYB_F64) { if (small) CYB_MID_LAUNCH(false, 8); else CYB_MID_LAUNCH(false, MID_MAX); }
    else { if (small) CYB_MID_LAUNCH(true, 8); else CYB_MID_LAUNCH(true, MID_MAX); }
#undef CYB_MID_LAUNCH
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // extern "C"
