// spline.cuh - scipy.interpolate.interp1d(kind="quadratic") = make_interp_spline(k=2) written out for one thread:
// not-a-knot knots, collocation matrix in 5-wide band storage, Gaussian elimination without pivoting (the matrix is totally
// positive), de Boor evaluation.  Shared by the A* trajectory initialiser (utils/math.py:57-65, astar.cu) and the path
// post-processor (ros/path_postprocessor.py:51-56, postprocess.cu); oracle/astar.py:quadratic_interp is the CPU restatement.
#pragma once

namespace nfopp {

// the three quadratic B-spline basis functions that are non-zero on [t[j], t[j+1]) (oracle/astar.py:_basis3)
__device__ __forceinline__ void quad_basis3(const double *t, double x, int j, double &n0, double &n1, double &n2) {
    const double left1 = x - t[j], right1 = t[j + 1] - x;
    const double b0 = right1 / (right1 + left1), b1 = left1 / (right1 + left1);
    const double left2 = x - t[j - 1], right2 = t[j + 2] - x;
    n0 = b0 * right1 / (right1 + left2);
    const double tmp = b0 * left2 / (right1 + left2);
    n1 = tmp + b1 * right2 / (right2 + left1);
    n2 = b1 * left1 / (right2 + left1);
}

// largest j in [from, n-1] with t[j] <= x (clamped); `from` = 2, or the previous result when x only grows
__device__ __forceinline__ int quad_interval(const double *t, int n, double x, int from) {
    int j = from;
    while (j < n - 1 && t[j + 1] <= x) ++j;
    return j;
}

// Fits the spline through (param[i], coef[NC * i + c]) for i < n (n >= 3, param strictly increasing): on return coef holds the
// B-spline coefficients and t[0 .. n+2] the knots.  band: n * 5 doubles of scratch.
template <int NC>
__device__ void quad_spline_fit(int n, const double *param, double *t, double *band, double *coef) {
    t[0] = t[1] = t[2] = param[0];
    for (int i = 1; i + 2 < n; ++i) t[2 + i] = (param[i + 1] + param[i]) / 2;      // interior midpoints without first / last
    t[n] = t[n + 1] = t[n + 2] = param[n - 1];
    for (int i = 0; i < n; ++i) {                                   // row i: B_{j-2..j}(param[i]) at band columns i-2 .. i+2
        for (int k = 0; k < 5; ++k) band[5 * i + k] = 0.0;
        const int j = quad_interval(t, n, param[i], 2);
        double n0, n1, n2;
        quad_basis3(t, param[i], j, n0, n1, n2);
        const int c0 = j - 2 - (i - 2);
        if (c0 >= 0 && c0 < 5) band[5 * i + c0] = n0;
        if (c0 + 1 >= 0 && c0 + 1 < 5) band[5 * i + c0 + 1] = n1;
        if (c0 + 2 >= 0 && c0 + 2 < 5) band[5 * i + c0 + 2] = n2;
    }
    for (int k = 0; k < n; ++k) {
        const double piv = band[5 * k + 2];
        for (int i = k + 1; i <= k + 2 && i < n; ++i) {
            const int ck = 2 - (i - k);                             // band column of matrix column k in row i
            const double f = band[5 * i + ck] / piv;
            if (f == 0.0) continue;
            for (int c = 0; c <= 2; ++c) {
                const int ci = ck + c;
                if (ci < 5) band[5 * i + ci] -= f * band[5 * k + 2 + c];
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) coef[NC * i + c] -= f * coef[NC * k + c];
        }
    }
    for (int i = n - 1; i >= 0; --i) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            double s = coef[NC * i + c];
            for (int d = 1; d <= 2 && i + d < n; ++d) s -= band[5 * i + 2 + d] * coef[NC * (i + d) + c];
            coef[NC * i + c] = s / band[5 * i + 2];
        }
    }
}

// value at x; j = interval hint from the previous (smaller) x, updated
template <int NC>
__device__ __forceinline__ void quad_spline_eval(const double *t, int n, const double *coef, double x, int &j, double (&y)[NC]) {
    j = quad_interval(t, n, x, j);
    double n0, n1, n2;
    quad_basis3(t, x, j, n0, n1, n2);
#pragma unroll
    for (int c = 0; c < NC; ++c) y[c] = n0 * coef[NC * (j - 2) + c] + n1 * coef[NC * (j - 1) + c] + n2 * coef[NC * j + c];
}

}  // namespace nfopp
