// Developer harness (not part of the product): times and checks libdba's device Cholesky + solve on
// random SPD systems of the reduced-system sizes (n = 6 P').  Build: tools/build_tools.sh
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../cut3r_slam_b200/csrc/dba_kernels.h"

namespace dba { thread_local int64_t g_launch_count = 0; void chol_trace_read(unsigned long long* out, int n); }
using namespace dba;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

static void cpu_solve(int n, const std::vector<double>& S, const std::vector<double>& v, std::vector<double>& x) {
  std::vector<double> L(S);
  for (int j = 0; j < n; ++j) {
    double d = L[j * n + j];
    for (int k = 0; k < j; ++k) d -= L[j * n + k] * L[j * n + k];
    d = std::sqrt(d);
    L[j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = L[i * n + j];
      for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
      L[i * n + j] = s / d;
    }
  }
  std::vector<double> y(n);
  for (int i = 0; i < n; ++i) { double s = v[i]; for (int k = 0; k < i; ++k) s -= L[i * n + k] * y[k]; y[i] = s / L[i * n + i]; }
  x.assign(n, 0);
  for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < n; ++k) s -= L[k * n + i] * x[k]; x[i] = s / L[i * n + i]; }
}

int main(int argc, char** argv) {
  std::vector<int> sizes = {60, 366, 756, 1530, 3060};
  double shift = 1.0;
  if (argc > 1) { sizes.clear(); for (int i = 1; i < argc; ++i) { if (argv[i][0] == 's') shift = atof(argv[i] + 1); else sizes.push_back(atoi(argv[i])); } }
  for (int n : sizes) {
    const int ld = (n + 1 + 3) & ~3;
    // SPD: S = G G^T / m + I with a banded-ish random G (cheap to build, well conditioned)
    std::vector<double> S(size_t(n) * n, 0.0), v(n);
    srand(1234 + n);
    const int m = 48;
    std::vector<double> G(size_t(n) * m);
    for (auto& g : G) g = (rand() / double(RAND_MAX)) - 0.5;
    for (int i = 0; i < n; ++i)
      for (int j = 0; j <= i; ++j) {
        double s = 0;
        for (int k = 0; k < m; ++k) s += G[size_t(i) * m + k] * G[size_t(j) * m + k];
        s = s / m + (i == j ? shift * (1.0 + 0.01 * i) : 0.0);
        S[size_t(i) * n + j] = s; S[size_t(j) * n + i] = s;
      }
    for (auto& e : v) e = (rand() / double(RAND_MAX)) - 0.5;
    std::vector<double> xref;
    cpu_solve(n, S, v, xref);
    std::vector<float> hA(size_t(n + 1) * ld, 0.f);
    for (int i = 0; i < n; ++i) for (int j = 0; j <= i; ++j) hA[size_t(i) * ld + j] = float(S[size_t(i) * n + j]);
    for (int j = 0; j < n; ++j) hA[size_t(n) * ld + j] = float(v[j]);

    Workspace ws{};
    float *A0;
    const size_t bytes = hA.size() * sizeof(float);
    CK(cudaMalloc(&A0, bytes + 4096)); CK(cudaMalloc(&ws.A, bytes + 4096)); CK(cudaMalloc(&ws.Lm, bytes + 4096));
    CK(cudaMalloc(&ws.Ldiag, cholesky_ldiag_floats(n) * 4 + 64)); CK(cudaMalloc(&ws.dx, n * 4 + 64)); CK(cudaMalloc(&ws.fail, 16));
    CK(cudaMemset(ws.fail, 0, 16)); CK(cudaMemset(ws.Lm, 0, bytes));
    ws.ldA = ld;
    CK(cudaMemcpy(A0, hA.data(), bytes, cudaMemcpyHostToDevice));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int max_ctas = cholesky_coop_max_ctas(n);
    for (int variant = 0; variant < 2; ++variant) {
      float best = 1e9f, sum = 0.f; const int reps = 20;
      for (int it = 0; it < reps + 3; ++it) {
        CK(cudaMemcpyAsync(ws.A, A0, bytes, cudaMemcpyDeviceToDevice, st));
        CK(cudaEventRecord(e0, st));
        if (variant == 0) CK(launch_cholesky_coop(ws, n, max_ctas, st)); else CK(launch_cholesky_solve(ws, n, st));
        CK(cudaEventRecord(e1, st));
        CK(cudaStreamSynchronize(st));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 3) { best = fminf(best, ms); sum += ms; }
      }
      std::vector<float> x(n);
      CK(cudaMemcpy(x.data(), ws.dx, n * 4, cudaMemcpyDeviceToHost));
#ifdef DBA_CHOL_TRACE
      if (variant == 0) {
        const int nkb = (n + 31) / 32;
        std::vector<unsigned long long> tr(size_t(nkb) * 16);
        chol_trace_read(tr.data(), nkb * 16);
        printf("   kernel (worker CTA 1, first to last stamp) %.1f us\n", double(tr[(nkb - 2) * 16 + 4] - tr[0]) * 1e-3);
        if (n == 756) for (int k = 0; k < nkb - 1; k += 3) printf("   step %2d: loads %llu trsm %llu update+store %llu barrier %llu | diag CTA %llu: loads %llu trsm %llu syrk %llu potrf %llu inv %llu\n", k, tr[k*16+1]-tr[k*16], tr[k*16+2]-tr[k*16+1], tr[k*16+3]-tr[k*16+2], tr[k*16+4]-tr[k*16+3], tr[k*16+7]-tr[k*16+6], tr[k*16+8]-tr[k*16+6], tr[k*16+9]-tr[k*16+8], tr[k*16+10]-tr[k*16+9], tr[k*16+11]-tr[k*16+10], tr[k*16+7]-tr[k*16+11]);
      }
#endif
      double err = 0, nrm = 0;
      for (int i = 0; i < n; ++i) { err = fmax(err, fabs(x[i] - xref[i])); nrm = fmax(nrm, fabs(xref[i])); }
      printf("n=%5d %-12s ctas=%3d  best %.1f us  mean %.1f us  max rel err %.2e\n", n, variant == 0 ? "fused-coop" : "multi-launch",
             max_ctas, best * 1e3, sum / reps * 1e3, err / nrm);
    }
    cudaFree(A0); cudaFree(ws.A); cudaFree(ws.Lm); cudaFree(ws.Ldiag); cudaFree(ws.dx); cudaFree(ws.fail);
  }
  return 0;
}
