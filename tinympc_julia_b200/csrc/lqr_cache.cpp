// See lqr_cache.hpp.  Reference behaviour: tinympc/TinyMPC/src/tinympc/tiny_api.cpp:107-108, 244-318.
#include "lqr_cache.hpp"

#include <cmath>
#include <utility>

namespace tmpc {

namespace {

template <class T>
Mat matmul_t(const Mat& X, const Mat& Y) {
    Mat Z(X.rows, Y.cols);
    for (int j = 0; j < Y.cols; ++j)
        for (int l = 0; l < X.cols; ++l) {
            const T y = static_cast<T>(Y(l, j));
            for (int i = 0; i < X.rows; ++i) Z(i, j) = static_cast<T>(static_cast<T>(Z(i, j)) + static_cast<T>(static_cast<T>(X(i, l)) * y));  // column AXPYs, ascending l per entry
        }
    return Z;
}

}  // namespace

Mat transpose(const Mat& X) {
    Mat T(X.cols, X.rows);
    for (int j = 0; j < X.cols; ++j)
        for (int i = 0; i < X.rows; ++i) T(j, i) = X(i, j);
    return T;
}

namespace {

template <class T>
Mat add_t(const Mat& X, const Mat& Y, double sy) {
    Mat Z(X.rows, X.cols);
    for (size_t k = 0; k < Z.a.size(); ++k) Z.a[k] = static_cast<T>(static_cast<T>(X.a[k]) + static_cast<T>(static_cast<T>(sy) * static_cast<T>(Y.a[k])));
    return Z;
}

template <class T>
bool invert_t(const Mat& X, Mat& Xinv) {
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
            lu(i, k) = static_cast<T>(static_cast<T>(lu(i, k)) / static_cast<T>(lu(k, k)));
            const T l = static_cast<T>(lu(i, k));
            for (int j = k + 1; j < n; ++j) lu(i, j) = static_cast<T>(static_cast<T>(lu(i, j)) - static_cast<T>(l * static_cast<T>(lu(k, j))));
        }
    }
    Xinv = Mat(n, n);
    for (int c = 0; c < n; ++c) {
        // forward substitution on the permuted unit vector, then back substitution
        for (int i = 0; i < n; ++i) {
            T s = (perm[i] == c) ? T(1) : T(0);
            for (int j = 0; j < i; ++j) s = static_cast<T>(s - static_cast<T>(static_cast<T>(lu(i, j)) * static_cast<T>(Xinv(j, c))));
            Xinv(i, c) = s;
        }
        for (int i = n - 1; i >= 0; --i) {
            T s = static_cast<T>(Xinv(i, c));
            for (int j = i + 1; j < n; ++j) s = static_cast<T>(s - static_cast<T>(static_cast<T>(lu(i, j)) * static_cast<T>(Xinv(j, c))));
            Xinv(i, c) = static_cast<T>(s / static_cast<T>(lu(i, i)));
        }
    }
    return true;
}

template <class T>
int compute_lqr_cache_t(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                        LqrCache& out) {
    auto matmul = [](const Mat& X, const Mat& Y) { return matmul_t<T>(X, Y); };
    auto add = [](const Mat& X, const Mat& Y, double sy = 1.0) { return add_t<T>(X, Y, sy); };
    auto invert = [](const Mat& X, Mat& Xi) { return invert_t<T>(X, Xi); };
    const int n = A.rows, m = B.cols;
    out.nx = n; out.nu = m; out.rho = rho;
    out.Qd.assign(n, 0.0);
    out.Rd.assign(m, 0.0);
    for (int i = 0; i < n; ++i) out.Qd[i] = static_cast<T>(static_cast<T>(Q(i, i)) + static_cast<T>(rho));   // what the loop multiplies Xref by
    for (int i = 0; i < m; ++i) out.Rd[i] = static_cast<T>(static_cast<T>(R(i, i)) + static_cast<T>(rho));

    // Riccati weights: diag(Qd)+rho*I and diag(Rd)+rho*I  (rho enters twice, as in the reference)
    Mat Q1(n, n), R1(m, m);
    for (int i = 0; i < n; ++i) Q1(i, i) = static_cast<T>(static_cast<T>(out.Qd[i]) + static_cast<T>(rho));
    for (int i = 0; i < m; ++i) R1(i, i) = static_cast<T>(static_cast<T>(out.Rd[i]) + static_cast<T>(rho));

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
        for (size_t k = 0; k < K.a.size(); ++k) delta = std::fmax(delta, std::fabs(static_cast<double>(static_cast<T>(static_cast<T>(K.a[k]) - static_cast<T>(Kprev.a[k])))));
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


}  // namespace

Mat matmul(const Mat& X, const Mat& Y) { return matmul_t<double>(X, Y); }
Mat add(const Mat& X, const Mat& Y, double sy) { return add_t<double>(X, Y, sy); }
bool invert(const Mat& X, Mat& Xinv) { return invert_t<double>(X, Xinv); }

int compute_lqr_cache(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                      LqrCache& out) {
    return compute_lqr_cache_t<double>(A, B, f, Q, R, rho, out);
}

// The float build of the reference (`typedef float tinytype`, types.hpp:15): inputs are rounded to float where
// src/bindings.cpp casts them (:50-55) and every operation of the recursion is rounded to float.  The result is a
// cache of float values held in doubles.  (The reference's float cache comes out of Eigen's float kernels, whose
// summation order differs; as for double, the two agree to the recursion's own 1e-5 stopping accuracy, and parity
// tests install the reference's cache through set_cache_terms.)
int compute_lqr_cache_f32(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                          LqrCache& out) {
    auto rnd = [](Mat m) { for (double& v : m.a) v = static_cast<double>(static_cast<float>(v)); return m; };
    return compute_lqr_cache_t<float>(rnd(A), rnd(B), rnd(f), rnd(Q), rnd(R),
                                      static_cast<double>(static_cast<float>(rho)), out);
}

}  // namespace tmpc
