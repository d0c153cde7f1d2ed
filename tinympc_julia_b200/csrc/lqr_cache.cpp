// See lqr_cache.hpp.  Reference behaviour: tinympc/TinyMPC/src/tinympc/tiny_api.cpp:107-108, 244-318.
#include "lqr_cache.hpp"

#include <cmath>
#include <utility>

namespace tmpc {

Mat matmul(const Mat& X, const Mat& Y) {
    Mat Z(X.rows, Y.cols);
    for (int j = 0; j < Y.cols; ++j)
        for (int l = 0; l < X.cols; ++l) {
            const double y = Y(l, j);
            for (int i = 0; i < X.rows; ++i) Z(i, j) += X(i, l) * y;  // column AXPYs, ascending l per entry
        }
    return Z;
}

Mat transpose(const Mat& X) {
    Mat T(X.cols, X.rows);
    for (int j = 0; j < X.cols; ++j)
        for (int i = 0; i < X.rows; ++i) T(j, i) = X(i, j);
    return T;
}

Mat add(const Mat& X, const Mat& Y, double sy) {
    Mat Z(X.rows, X.cols);
    for (size_t k = 0; k < Z.a.size(); ++k) Z.a[k] = X.a[k] + sy * Y.a[k];
    return Z;
}

bool invert(const Mat& X, Mat& Xinv) {
    const int n = X.rows;
    Mat lu = X;
    std::vector<int> perm(n);
    for (int i = 0; i < n; ++i) perm[i] = i;
    for (int k = 0; k < n; ++k) {
        int piv = k;
        double best = std::fabs(lu(k, k));
        for (int i = k + 1; i < n; ++i)
            if (std::fabs(lu(i, k)) > best) { best = std::fabs(lu(i, k)); piv = i; }
        if (!(best > 0.0) || !std::isfinite(best)) return false;
        if (piv != k) {
            for (int j = 0; j < n; ++j) std::swap(lu(k, j), lu(piv, j));
            std::swap(perm[k], perm[piv]);
        }
        for (int i = k + 1; i < n; ++i) {
            lu(i, k) /= lu(k, k);
            const double l = lu(i, k);
            for (int j = k + 1; j < n; ++j) lu(i, j) -= l * lu(k, j);
        }
    }
    Xinv = Mat(n, n);
    for (int c = 0; c < n; ++c) {
        // forward substitution on the permuted unit vector, then back substitution
        for (int i = 0; i < n; ++i) {
            double s = (perm[i] == c) ? 1.0 : 0.0;
            for (int j = 0; j < i; ++j) s -= lu(i, j) * Xinv(j, c);
            Xinv(i, c) = s;
        }
        for (int i = n - 1; i >= 0; --i) {
            double s = Xinv(i, c);
            for (int j = i + 1; j < n; ++j) s -= lu(i, j) * Xinv(j, c);
            Xinv(i, c) = s / lu(i, i);
        }
    }
    return true;
}

int compute_lqr_cache(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                      LqrCache& out) {
    const int n = A.rows, m = B.cols;
    out.nx = n; out.nu = m; out.rho = rho;
    out.Qd.assign(n, 0.0);
    out.Rd.assign(m, 0.0);
    for (int i = 0; i < n; ++i) out.Qd[i] = Q(i, i) + rho;   // what the loop multiplies Xref by
    for (int i = 0; i < m; ++i) out.Rd[i] = R(i, i) + rho;

    // Riccati weights: diag(Qd)+rho*I and diag(Rd)+rho*I  (rho enters twice, as in the reference)
    Mat Q1(n, n), R1(m, m);
    for (int i = 0; i < n; ++i) Q1(i, i) = out.Qd[i] + rho;
    for (int i = 0; i < m; ++i) R1(i, i) = out.Rd[i] + rho;

    const Mat At = transpose(A), Bt = transpose(B);
    Mat P(n, n), K(m, n), Kprev(m, n), Pnext(n, n);
    for (int i = 0; i < n; ++i) P(i, i) = rho;

    int sweeps = 0;
    for (int it = 0; it < 1000; ++it) {
        sweeps = it + 1;
        Mat G = add(R1, matmul(matmul(Bt, P), B));
        Mat Gi;
        if (!invert(G, Gi)) return 1;
        K = matmul(matmul(matmul(Gi, Bt), P), A);                 // ((G^-1 B') P) A
        Mat AmBK = add(A, matmul(B, K), -1.0);
        Pnext = add(Q1, matmul(matmul(At, P), AmBK));              // Q1 + (A' P)(A - B K)
        double delta = 0.0;
        for (size_t k = 0; k < K.a.size(); ++k) delta = std::fmax(delta, std::fabs(K.a[k] - Kprev.a[k]));
        if (delta < 1e-5) break;      // Pnext of this sweep is the one kept
        Kprev = K;
        P = Pnext;
    }
    out.riccati_sweeps = sweeps;
    out.Kinf = K;
    out.Pinf = Pnext;

    Mat G = add(R1, matmul(matmul(Bt, out.Pinf), B));
    if (!invert(G, out.Quu_inv)) return 1;
    out.AmBKt = transpose(add(A, matmul(B, out.Kinf), -1.0));
    out.APf = matmul(matmul(out.AmBKt, out.Pinf), f);
    out.BPf = matmul(matmul(Bt, out.Pinf), f);
    return 0;
}

}  // namespace tmpc
