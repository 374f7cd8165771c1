// minsnap.cu — batched linear min-snap solve: waypoint sets + segment times in, polynomial
// coefficients out in the layout the fused sweep K3b reads ([trajectory][segment][dim][coefficient]).
//
// Replaces, for the vertex pattern poly_traj::generateTrajectory builds (external/poly_traj/src/
// trajectory_generator.cpp:12-58: start = position / velocity / acceleration given with jerk = snap = 0, interior
// vertices = position only, end = position with all derivatives 0), the reference's
// PolynomialOptimization<10>::setupFromVertices + solveLinear
// (external/poly_traj/include/mav_trajectory_generation/impl/polynomial_optimization_linear_impl.h:56-380), which
// builds sparse matrices and runs Eigen's SparseQR once per trajectory on the CPU.
//
// Formulation (Richter et al., the reference's [1]): per segment H = A^-T Q A^-1 in end-point-derivative
// coordinates, R = C^T H C over (vertex, derivative) variables, free derivatives dp = -Rpp^-1 Rpf df.  A segment
// of duration T is the unit segment under u_r = T^r d_r and b_j = T^j c_j with the cost scaled by T^-7, so
// H(T) = T^-7 D H1 D with D = diag(T^r) and the constant H1 = A1^-T Q1 A1^-1 (minsnap_constants.h, exact
// rational arithmetic rounded once): nothing of powers of T is ever inverted.  Rpp is symmetric positive
// definite of order 4 (K - 1) and block tridiagonal (a free derivative only meets those of the neighbouring
// vertices), stored as a lower band of width 8: one warp per trajectory assembles it in shared memory (fixed lane -> row
// assignment, no atomics: results are deterministic), factors it (Cholesky), solves the three dimensions and
// maps the end-point derivatives of every segment back to coefficients.
//
// Floating point: binary64, no FMA contraction (the TU is built with -fmad=false like the rest of the library).
// This is a floating-point kernel: parity is by tolerance (1e-9 m along every segment, tests/test_minsnap.py) against
// the reference's own PolynomialOptimization<10> code compiled over stand-in linear algebra (oracle/_ref, golden
// vectors) and against oracle/minsnap_np.py, a numpy restatement of the same unscaled formulation.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>

#include "kernels.h"
#include "minsnap_constants.h"

namespace epp {

namespace {

constexpr int kMsN = 10;       // coefficients per polynomial
constexpr int kMsHalf = 5;     // end-point derivatives per vertex: position .. snap
constexpr int kMsMaxSeg = 128; // segments per trajectory supported by the shared-memory layout
constexpr int kMsBand = 8;     // Rpp is block tridiagonal in 4 x 4 blocks: |i - j| <= 7; lower band storage [row][row - col]

__constant__ double c_A1inv[kMsN][kMsN];
__constant__ double c_H1[kMsN][kMsN];

extern __shared__ __align__(16) unsigned char ms_smem[];

// free-variable index of (vertex v, derivative r) or -1 when the variable is fixed
__device__ __forceinline__ int ms_free_index(int v, int r, int K) { return (v >= 1 && v < K && r >= 1) ? 4 * (v - 1) + (r - 1) : -1; }

// value of the fixed variable (vertex v, derivative r) in dimension d
__device__ __forceinline__ double ms_fixed_value(const double *wp, const double *va, int v, int r, int d, int K)
{
    if (r == 0) return wp[3 * v + d];
    if (v == 0 && va) return r == 1 ? va[d] : (r == 2 ? va[3 + d] : 0.0);
    return 0.0;
}

__global__ void __launch_bounds__(32) k_minsnap(const double *__restrict__ waypoints, const double *__restrict__ durations,
                                                const double *__restrict__ start_vel_acc, size_t n_traj, int K,
                                                double *__restrict__ coeffs)
{
    const size_t t = blockIdx.x;
    if (t >= n_traj) return;
    const int lane = threadIdx.x;
    const int nP = 4 * (K - 1);
    // shared: Tp[K][10] powers of T | wp[(K+1)*3] | Rpp lower band [nP][8] | rhs[3][nP]
    double *Tp = reinterpret_cast<double *>(ms_smem);
    double *wp = Tp + K * kMsN;
    double *R = wp + 3 * (K + 1);
    double *rhs = R + nP * kMsBand;
    const double *va = start_vel_acc ? start_vel_acc + 6 * t : nullptr;

    for (int i = lane; i < 3 * (K + 1); i += 32) wp[i] = waypoints[t * 3 * (size_t)(K + 1) + i];
    for (int k = lane; k < K; k += 32) {
        const double T = durations[t * (size_t)K + k];
        double p = 1.0;
        for (int j = 0; j < kMsN; ++j) {
            Tp[k * kMsN + j] = p;
            p = p * T;
        }
    }
    for (int i = lane; i < nP * kMsBand + 3 * nP; i += 32) R[i] = 0.0; // Rpp and rhs are contiguous
    __syncwarp();

    // ---- assemble Rpp and rhs = -Rpf df: lane = local row of the segment's 10 x 10 block, segments in turn
    for (int k = 0; k < K; ++k) {
        if (lane < kMsN) {
            const int row = lane, vr = k + row / kMsHalf, rr = row % kMsHalf;
            const int fi = ms_free_index(vr, rr, K);
            if (fi >= 0) {
                const double *tp = Tp + k * kMsN;
                const double inv_t7 = 1.0 / tp[7];
                const double srow = tp[rr] * inv_t7;
                double b0 = 0.0, b1 = 0.0, b2 = 0.0;
                for (int col = 0; col < kMsN; ++col) {
                    const int vc = k + col / kMsHalf, rc = col % kMsHalf;
                    const double h = c_H1[row][col] * srow * tp[rc];
                    const int fj = ms_free_index(vc, rc, K);
                    if (fj >= 0) {
                        if (fj <= fi) R[fi * kMsBand + (fi - fj)] += h; // lower band only (H1 is symmetric)
                    } else {
                        b0 += h * ms_fixed_value(wp, va, vc, rc, 0, K);
                        b1 += h * ms_fixed_value(wp, va, vc, rc, 1, K);
                        b2 += h * ms_fixed_value(wp, va, vc, rc, 2, K);
                    }
                }
                rhs[fi] -= b0;
                rhs[nP + fi] -= b1;
                rhs[2 * nP + fi] -= b2;
            }
        }
        __syncwarp(); // the next segment shares its start vertex's rows with this one's end vertex
    }

    // ---- banded Cholesky Rpp = L L^T in place, right-looking: column j touches rows j+1 .. j+7 only
#define MS_L(i, c) R[(i) * kMsBand + ((i) - (c))]
    for (int j = 0; j < nP; ++j) {
        if (lane == 0) MS_L(j, j) = sqrt(MS_L(j, j));
        __syncwarp();
        const double ljj = MS_L(j, j);
        const int last = (j + kMsBand - 1) < (nP - 1) ? (j + kMsBand - 1) : (nP - 1);
        if (lane >= 1 && j + lane <= last) MS_L(j + lane, j) = MS_L(j + lane, j) / ljj;
        __syncwarp();
        // trailing update: lanes over the (i, c) pairs j < c <= i <= last (at most 28)
        const int rows = last - j;
        if (lane < rows * (rows + 1) / 2) {
            int a = 0, rem = lane; // a-th row below j, rem-th column in it
            while (rem > a) {
                rem -= a + 1;
                ++a;
            }
            const int i = j + 1 + a, c = j + 1 + rem;
            MS_L(i, c) -= MS_L(i, j) * MS_L(c, j);
        }
        __syncwarp();
    }
    // ---- L y = b, L^T x = y: lanes 0..2 take one dimension each
    if (lane < 3) {
        double *b = rhs + lane * nP;
        for (int i = 0; i < nP; ++i) {
            double s = b[i];
            for (int c = (i - (kMsBand - 1)) > 0 ? (i - (kMsBand - 1)) : 0; c < i; ++c) s -= MS_L(i, c) * b[c];
            b[i] = s / MS_L(i, i);
        }
        for (int i = nP - 1; i >= 0; --i) {
            double s = b[i];
            const int last = (i + kMsBand - 1) < (nP - 1) ? (i + kMsBand - 1) : (nP - 1);
            for (int c = i + 1; c <= last; ++c) s -= MS_L(c, i) * b[c];
            b[i] = s / MS_L(i, i);
        }
    }
#undef MS_L
    __syncwarp();

    // ---- coefficients: lanes over (segment, dimension); c_j = (A1inv (D d))_j / T^j
    for (int kd = lane; kd < 3 * K; kd += 32) {
        const int k = kd / 3, d = kd % 3;
        const double *tp = Tp + k * kMsN;
        double u[kMsN];
#pragma unroll
        for (int row = 0; row < kMsN; ++row) {
            const int v = k + row / kMsHalf, r = row % kMsHalf;
            const int fi = ms_free_index(v, r, K);
            const double val = fi >= 0 ? rhs[d * nP + fi] : ms_fixed_value(wp, va, v, r, d, K);
            u[row] = tp[r] * val;
        }
        double *out = coeffs + ((t * (size_t)K + k) * 3 + d) * kMsN;
#pragma unroll
        for (int j = 0; j < kMsN; ++j) {
            double s = 0.0;
#pragma unroll
            for (int row = 0; row < kMsN; ++row) s += c_A1inv[j][row] * u[row];
            out[j] = s / tp[j];
        }
    }
}

// ---- trajectory matrix: the sampling half of poly_traj::generateTrajectory ---------------------------------
//
// trajectory_generator.cpp:60-97: three evaluateRange passes (position, velocity, acceleration) over
// [getMinTime(), getMaxTime()] at the sampling interval, written as rows
// [px vx ax  py vy ay  pz vz az  t + startTimeOffset].  One warp per trajectory: lane 0 advances the reference's
// sequential sampler (trajectory.cpp:81-141, running sums) by 32 samples, then the 32 lanes evaluate one sample
// each: Polynomial::evaluate(t, derivative) (polynomial.h:136-150) — Horner from the top coefficient with the
// base coefficients j! / (j - derivative)! (polynomial.cpp:145-160), every operation individually rounded, so
// the matrix is bit-equal to the reference's.  Output is column-major rows x 10 with leading dimension ld
// (Eigen's MatrixXd layout), one matrix per trajectory.
__device__ __forceinline__ double ms_eval(const double *__restrict__ c, int n, double t, int derivative)
{
    auto bc = [&](int j) -> double { // base_coefficients_(derivative, j)
        double v = 1.0;
        for (int r = 0; r < derivative; ++r) v *= (double)(j - r);
        return v;
    };
    double r = __dmul_rn(bc(n - 1), c[n - 1]);
    for (int j = n - 2; j >= derivative; --j) {
        r = __dmul_rn(r, t);
        r = __dadd_rn(r, __dmul_rn(bc(j), c[j]));
    }
    return r;
}

__global__ void __launch_bounds__(32) k_sample_poly(const double *__restrict__ coeffs, const double *__restrict__ durations,
                                                    const double *__restrict__ dts, const double *__restrict__ time_offsets,
                                                    size_t n_traj, int K, int N, size_t ld, double *__restrict__ out,
                                                    uint32_t *__restrict__ n_rows)
{
    __shared__ double s_t[32], s_acc[32];
    __shared__ int s_seg[32];
    const size_t tr = blockIdx.x;
    if (tr >= n_traj) return;
    const int lane = threadIdx.x;
    const double *dur = durations + tr * (size_t)K;
    const double dt = dts[tr], off = time_offsets ? time_offsets[tr] : 0.0;
    double *m = out + tr * 10 * ld;
    // sampler state (lane 0 is the only one that advances it)
    double t_end = 0.0, acc = 0.0, tin = 0.0, seg_T = 0.0;
    int seg = 0;
    bool done = false;
    if (lane == 0) {
        for (int k = 0; k < K; ++k) t_end = __dadd_rn(t_end, dur[k]); // trajectory.h:64-71
        double a = 0.0;
        int k = 0;
        for (k = 0; k < K; ++k) { // start-segment search with t_start = 0 (trajectory.cpp:99-110)
            a = __dadd_rn(a, dur[k]);
            if (a > 0.0) break;
        }
        if (0.0 > a || k >= K) {
            done = true;
        } else {
            seg = k;
            seg_T = dur[seg];
            acc = __dsub_rn(a, seg_T);
            tin = __dsub_rn(0.0, acc);
        }
    }
    size_t rows = 0;
    for (;;) {
        int cnt = 0;
        if (lane == 0) {
            while (!done && cnt < 32) {
                if (!(acc < t_end)) {
                    done = true;
                    break;
                }
                if (tin > seg_T) {
                    tin = __dsub_rn(tin, seg_T);
                    ++seg;
                    if (seg >= K) {
                        done = true;
                        break;
                    }
                    seg_T = dur[seg];
                    continue;
                }
                s_t[cnt] = tin;
                s_acc[cnt] = acc;
                s_seg[cnt] = seg;
                ++cnt;
                tin = __dadd_rn(tin, dt);
                acc = __dadd_rn(acc, dt);
            }
        }
        cnt = __shfl_sync(0xffffffffu, cnt, 0);
        if (cnt == 0) break;
        __syncwarp();
        const size_t row = rows + (size_t)lane;
        if (lane < cnt && row < ld) {
            const double t = s_t[lane];
            const double *c = coeffs + ((tr * (size_t)K + (size_t)s_seg[lane]) * 3) * (size_t)N;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                m[(size_t)(3 * d) * ld + row] = ms_eval(c + d * N, N, t, 0);
                m[(size_t)(3 * d + 1) * ld + row] = N > 1 ? ms_eval(c + d * N, N, t, 1) : 0.0;
                m[(size_t)(3 * d + 2) * ld + row] = N > 2 ? ms_eval(c + d * N, N, t, 2) : 0.0;
            }
            m[9 * ld + row] = __dadd_rn(s_acc[lane], off);
        }
        rows += (size_t)cnt;
        __syncwarp();
    }
    if (lane == 0 && n_rows) n_rows[tr] = (uint32_t)rows;
}

bool g_ms_constants_ready[64] = {};
std::mutex g_ms_mu;

} // namespace

cudaError_t launch_minsnap(const double *d_waypoints, const double *d_durations, const double *d_start_vel_acc,
                           size_t n_traj, int n_segments, double *d_coeffs, cudaStream_t stream)
{
    if (n_traj == 0) return cudaSuccess;
    if (n_segments < 1 || n_segments > kMsMaxSeg || n_traj > 0x7fffffffull) return cudaErrorInvalidValue;
    int device = 0;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return err;
    std::lock_guard<std::mutex> lk(g_ms_mu); // first launch per device uploads the two constant matrices
    if (device >= 64) return cudaErrorInvalidDevice;
    if (!g_ms_constants_ready[device]) {
        err = cudaMemcpyToSymbol(c_A1inv, kMinsnapA1inv, sizeof(kMinsnapA1inv));
        if (err == cudaSuccess) err = cudaMemcpyToSymbol(c_H1, kMinsnapH1, sizeof(kMinsnapH1));
        if (err != cudaSuccess) return err;
        g_ms_constants_ready[device] = true;
    }
    const int K = n_segments, nP = 4 * (K - 1);
    const size_t smem = sizeof(double) * ((size_t)K * kMsN + 3 * (K + 1) + (size_t)nP * kMsBand + 3 * nP);
    if (smem > 40 * 1024) {
        err = cudaFuncSetAttribute(k_minsnap, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
    }
    k_minsnap<<<(unsigned)n_traj, 32, smem, stream>>>(d_waypoints, d_durations, d_start_vel_acc, n_traj, K, d_coeffs);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_sample_poly(const double *d_coeffs, const double *d_durations, const double *d_dt,
                               const double *d_time_offsets, size_t n_traj, int n_segments, int n_coeffs, size_t ld,
                               double *d_out, uint32_t *d_n_rows, cudaStream_t stream)
{
    if (n_traj == 0) return cudaSuccess;
    if (n_segments < 1 || n_coeffs < 1 || ld == 0 || n_traj > 0x7fffffffull) return cudaErrorInvalidValue;
    k_sample_poly<<<(unsigned)n_traj, 32, 0, stream>>>(d_coeffs, d_durations, d_dt, d_time_offsets, n_traj, n_segments,
                                                      n_coeffs, ld, d_out, d_n_rows);
    g_launches++;
    return cudaGetLastError();
}

} // namespace epp
