// Host-side infinite-horizon LQR cache for the ADMM solve loop.
//
// Produces every constant the device loop reads: Kinf, Pinf, Quu_inv, AmBKt, APf, BPf plus the
// rho-augmented cost diagonals Qd, Rd.  Runs once per problem on the host (it is O(1000 * n^3)
// flops on 12x12 matrices); the solve kernels only consume the result.
//
// Behaviour follows the reference's tiny_setup / tiny_precompute_and_set_cache
// (reference: tinympc/TinyMPC/src/tinympc/tiny_api.cpp:107-108, 244-318), including its quirks:
//   * only the DIAGONALS of Q and R are used, and rho is added twice on the way into the
//     Riccati recursion (once at :107-108, again at :254-255);
//   * the recursion starts from P = rho*I, runs at most 1000 sweeps and stops when
//     max|Kinf - Kinf_prev| < 1e-5.
#pragma once
#include <cstddef>
#include <vector>

namespace tmpc {

// Column-major dense matrix, just enough for the setup math.
struct Mat {
    int rows = 0, cols = 0;
    std::vector<double> a;
    Mat() = default;
    Mat(int r, int c) : rows(r), cols(c), a(static_cast<size_t>(r) * c, 0.0) {}
    Mat(int r, int c, const double* src) : rows(r), cols(c), a(src, src + static_cast<size_t>(r) * c) {}
    double& operator()(int i, int j) { return a[static_cast<size_t>(j) * rows + i]; }
    double operator()(int i, int j) const { return a[static_cast<size_t>(j) * rows + i]; }
    const double* data() const { return a.data(); }
    double* data() { return a.data(); }
    size_t size() const { return a.size(); }
};

Mat matmul(const Mat& X, const Mat& Y);
Mat transpose(const Mat& X);
Mat add(const Mat& X, const Mat& Y, double sy = 1.0);  // X + sy*Y
bool invert(const Mat& X, Mat& Xinv);                  // LU with partial pivoting; false if singular

struct LqrCache {
    int nx = 0, nu = 0;
    double rho = 0.0;
    Mat Kinf;     // nu x nx
    Mat Pinf;     // nx x nx
    Mat Quu_inv;  // nu x nu
    Mat AmBKt;    // nx x nx
    Mat APf;      // nx x 1
    Mat BPf;      // nu x 1
    std::vector<double> Qd, Rd;  // diag(Q)+rho, diag(R)+rho : the only part of Q, R the loop sees
    int riccati_sweeps = 0;
};

// 0 on success, 1 when a matrix that must be inverted is singular.
int compute_lqr_cache(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                      LqrCache& out);

// The same recursion with every input and operation rounded to float (the reference built with
// `typedef float tinytype`); the values come back as doubles.
int compute_lqr_cache_f32(const Mat& A, const Mat& B, const Mat& f, const Mat& Q, const Mat& R, double rho,
                          LqrCache& out);

}  // namespace tmpc
