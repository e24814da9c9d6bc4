// gpu_library.cu -- pybind11 drop-in for the reference module of the same name.
//
// Keeps the surface of /root/reference/gpu_library.cu:85-87, 236-239:
//   gpu_library.linsys_solve(G_row, G_col, G_val, C_row, C_col, C_val, g_val, c_val, input_lambda,
//                            testiters, exit_tol, max_iters, warm_start, rho) -> (lambda, dz)
// with the compile-time knobs STATE_SIZE / CONTROL_SIZE / KNOT_POINTS (CMakeLists.txt:18).
// Everything numerical happens behind the C ABI (include/gato_b200.h).  Differences from the
// reference, all at the boundary (SURVEY.md 8b "EXT"): argument sizes are validated (ValueError
// instead of undefined behaviour), CUDA failures raise RuntimeError instead of exit(), the
// iteration count is also returned by linsys_solve_info(), the plan (pattern analysis + device
// workspace) persists between calls with the same sparsity pattern.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gato_b200.h"

#if !defined(STATE_SIZE) || !defined(CONTROL_SIZE) || !defined(KNOT_POINTS)
#error "build with -DSTATE_SIZE=.. -DCONTROL_SIZE=.. -DKNOT_POINTS=.."
#endif

namespace py = pybind11;

namespace {

constexpr int kS = STATE_SIZE, kC = CONTROL_SIZE, kK = KNOT_POINTS;
constexpr int kNz = (kS + kC) * kK - kC, kNl = kS * kK;

struct Cache {
    gato_plan* plan = nullptr;
    std::vector<int> G_row, G_col, C_row, C_col;
    int last_iters = -1;
    float first_ms = 0.f, avg_ms = 0.f;
    bool true_warm = false;
    ~Cache() { /* the CUDA context may already be gone at interpreter exit: leak the plan */ }
} g_cache;

// Threading.  The reference holds the GIL for the whole call (SURVEY 8b "Threading"), which serialises its callers.
// Here the GIL is released around the device work, so the process-global plan cache and the plan's staging buffers
// are guarded by g_mu instead: a call takes it AFTER dropping the GIL (never the other way round) and holds it
// across plan lookup / re-creation and the solve; errors are carried out of the locked region and raised with the
// GIL held.
std::mutex g_mu;

struct Err {
    int rc = GATO_OK;
    std::string msg;
    void set(int rc_) { if (rc_ != GATO_OK && rc == GATO_OK) { rc = rc_; msg = gato_last_error(); } }
    void raise() const {
        if (rc == GATO_OK) return;
        if (rc == GATO_EINVAL || rc == GATO_EPATTERN) throw py::value_error(msg);
        throw std::runtime_error("gato_b200: " + msg + " (code " + std::to_string(rc) + ")");
    }
};

void check(int rc) {
    Err e;
    e.set(rc);
    e.raise();
}

template <class F>
void locked_nogil(F&& f) {
    py::gil_scoped_release nogil;
    std::lock_guard<std::mutex> lk(g_mu);
    f();
}

bool quiet() {
    const char* q = std::getenv("GATO_B200_QUIET");
    return q && q[0] && q[0] != '0';
}

// call with g_mu held; nullptr + err on failure
gato_plan* plan_for(const std::vector<int>& G_row, const std::vector<int>& G_col,
                    const std::vector<int>& C_row, const std::vector<int>& C_col, Err* err) {
    Cache& c = g_cache;
    if (c.plan && c.G_row == G_row && c.G_col == G_col && c.C_row == C_row && c.C_col == C_col)
        return c.plan;
    if (c.plan) { gato_plan_destroy(c.plan); c.plan = nullptr; }
    int dev = 0;
    if (const char* d = std::getenv("GATO_B200_DEVICE")) dev = std::atoi(d);
    err->set(gato_plan_create(kS, kC, kK, G_row.data(), G_col.data(), (int)G_col.size(), C_row.data(),
                              C_col.data(), (int)C_col.size(), dev, 1, &c.plan));
    if (err->rc) { c.plan = nullptr; return nullptr; }
    c.G_row = G_row; c.G_col = G_col; c.C_row = C_row; c.C_col = C_col;
    return c.plan;
}

void check_sizes(size_t G_row, size_t G_col, size_t G_val, size_t C_row, size_t C_col, size_t C_val,
                 size_t g, size_t c, size_t lam) {
    auto bad = [](const char* name, size_t got, size_t want) {
        throw py::value_error(std::string(name) + " has " + std::to_string(got) + " entries, expected " +
                              std::to_string(want) + " for STATE_SIZE=" + std::to_string(kS) +
                              " CONTROL_SIZE=" + std::to_string(kC) + " KNOT_POINTS=" + std::to_string(kK));
    };
    if (G_row != (size_t)kNz + 1) bad("G_row", G_row, kNz + 1);
    if (C_row != (size_t)kNl + 1) bad("C_row", C_row, kNl + 1);
    if (G_col != G_val) bad("G_val", G_val, G_col);
    if (C_col != C_val) bad("C_val", C_val, C_col);
    if (g != (size_t)kNz) bad("g_val", g, kNz);
    if (c != (size_t)kNl) bad("c_val", c, kNl);
    if (lam < (size_t)kNl) bad("input_lambda", lam, kNl);
}

py::tuple linsys_solve(std::vector<int> G_row, std::vector<int> G_col, std::vector<float> G_val,
                       std::vector<int> C_row, std::vector<int> C_col, std::vector<float> C_val,
                       std::vector<float> g_val, std::vector<float> c_val,
                       std::vector<float> input_lambda, int testiters, float exit_tol, int max_iters,
                       bool warm_start, float rho) {
    check_sizes(G_row.size(), G_col.size(), G_val.size(), C_row.size(), C_col.size(), C_val.size(),
                g_val.size(), c_val.size(), input_lambda.size());
    if (testiters < 1) throw py::value_error("testiters must be >= 1");
    std::vector<float> lambda(input_lambda.begin(), input_lambda.begin() + kNl), lam_out(kNl), dz(kNz);
    double total = 0.0;
    int first_iters = 0;
    float first_ms = 0.f, avg_ms = 0.f;
    Err err;
    locked_nogil([&] {
        gato_plan* plan = plan_for(G_row, G_col, C_row, C_col, &err);
        if (!plan) return;
        const int warm = !warm_start ? GATO_WARM_OFF
                                     : (g_cache.true_warm ? GATO_WARM_TRUE : GATO_WARM_REFERENCE);
        for (int i = 0; i < testiters; ++i) {   // the reference repeats the whole solve (gpu_library.cu:169-192)
            int iters = 0;
            float ms = 0.f;
            err.set(gato_solve(plan, G_val.data(), C_val.data(), g_val.data(), c_val.data(), lambda.data(),
                               lam_out.data(), dz.data(), exit_tol, max_iters, warm, rho, &iters, &ms));
            if (err.rc) return;
            total += ms;
            if (i == 0) { first_iters = iters; first_ms = ms; }
            g_cache.last_iters = iters;
        }
        g_cache.first_ms = first_ms;
        g_cache.avg_ms = avg_ms = (float)(total / testiters);
    });
    err.raise();
    if (!quiet()) {   // same two summary lines as gpu_library.cu:190,198 (log-scraping callers)
        std::printf("first run PCG terminated in %d iterations, time:  %f\n", first_iters, first_ms);
        std::printf("avg time: %f\n", avg_ms);
        std::fflush(stdout);
    }
    py::list l, d;
    for (int i = 0; i < kNl; ++i) l.append(lam_out[i]);
    for (int i = 0; i < kNz; ++i) d.append(dz[i]);
    return py::make_tuple(l, d);
}

py::dict linsys_solve_info() {
    int iters = -1, cl = 0, kpc = 0;
    float first_ms = 0.f, avg_ms = 0.f;
    bool have = false;
    locked_nogil([&] {
        iters = g_cache.last_iters; first_ms = g_cache.first_ms; avg_ms = g_cache.avg_ms;
        have = g_cache.plan && gato_plan_get_cluster(g_cache.plan, &cl, &kpc) == GATO_OK;
    });
    py::dict d;
    d["iters"] = iters;
    d["first_ms"] = first_ms;
    d["avg_ms"] = avg_ms;
    if (have) {
        d["cluster_ctas"] = cl;
        d["knots_per_cta"] = kpc;
    }
    return d;
}

using farr = py::array_t<float, py::array::c_style | py::array::forcecast>;
using iarr = py::array_t<int, py::array::c_style | py::array::forcecast>;

// EXT: batch of systems over one pattern, NumPy in / NumPy out, no list boxing.
py::tuple linsys_solve_batched(iarr G_row, iarr G_col, farr G_val, iarr C_row, iarr C_col, farr C_val,
                               farr g_val, farr c_val, float exit_tol, int max_iters, float rho) {
    if (G_val.ndim() != 2 || C_val.ndim() != 2 || g_val.ndim() != 2 || c_val.ndim() != 2)
        throw py::value_error("G_val, C_val, g_val, c_val must be [nsys, n] arrays");
    const int nsys = (int)g_val.shape(0);
    check_sizes(G_row.size(), G_col.size(), G_val.shape(1), C_row.size(), C_col.size(), C_val.shape(1),
                g_val.shape(1), c_val.shape(1), kNl);
    auto rows_ok = [&](const farr& a) { return a.shape(0) == nsys || a.shape(0) == 1; };
    if (!rows_ok(G_val) || !rows_ok(C_val) || !rows_ok(c_val))
        throw py::value_error("leading dimensions must be nsys or 1 (shared)");
    std::vector<int> gr(G_row.data(), G_row.data() + G_row.size()), gc(G_col.data(), G_col.data() + G_col.size());
    std::vector<int> cr(C_row.data(), C_row.data() + C_row.size()), cc(C_col.data(), C_col.data() + C_col.size());
    py::array_t<float> lam({nsys, kNl}), dz({nsys, kNz});
    py::array_t<int> iters(nsys);
    auto ld = [&](const farr& a) -> long long { return a.shape(0) == 1 && nsys > 1 ? 0 : (long long)a.shape(1); };
    const long long lG = ld(G_val), lC = ld(C_val), lg = ld(g_val), lc = ld(c_val);
    const float *pG = G_val.data(), *pC = C_val.data(), *pg = g_val.data(), *pc = c_val.data();
    float* pl_ = lam.mutable_data(); float* pd = dz.mutable_data(); int* pi = iters.mutable_data();
    Err err;
    locked_nogil([&] {
        gato_plan* plan = plan_for(gr, gc, cr, cc, &err);
        if (!plan) return;
        float ms = 0.f;
        err.set(gato_solve_batched(plan, nsys, pG, lG, pC, lC, pg, lg, pc, lc, nullptr, 0, pl_, pd, pi, exit_tol,
                                   max_iters, GATO_WARM_OFF, rho, &ms));
        g_cache.avg_ms = g_cache.first_ms = ms;
    });
    err.raise();
    return py::make_tuple(lam, dz, iters);
}

// EXT: block input, no CSR.  G_dense [nsys, (S^2+C^2)K-C^2] and C_dense
// [nsys, (S^2+SC)(K-1)] are the reference's dense layouts (include/gato_defines.h:36-37); a leading
// dimension of 1 shares the array between all systems.
gato_plan* g_block_plan = nullptr;

py::tuple linsys_solve_blocks(farr G_dense, farr C_dense, farr g_val, farr c_val, float exit_tol, int max_iters,
                              float rho) {
    constexpr int nG = (kS * kS + kC * kC) * kK - kC * kC, nC = (kS * kS + kS * kC) * (kK - 1);
    if (G_dense.ndim() != 2 || C_dense.ndim() != 2 || g_val.ndim() != 2 || c_val.ndim() != 2)
        throw py::value_error("G_dense, C_dense, g_val, c_val must be [nsys, n] arrays");
    const int nsys = (int)g_val.shape(0);
    if (G_dense.shape(1) != nG || C_dense.shape(1) != nC || g_val.shape(1) != kNz || c_val.shape(1) != kNl)
        throw py::value_error("expected row lengths " + std::to_string(nG) + ", " + std::to_string(nC) + ", " +
                              std::to_string(kNz) + ", " + std::to_string(kNl));
    auto rows_ok = [&](const farr& a) { return a.shape(0) == nsys || a.shape(0) == 1; };
    if (!rows_ok(G_dense) || !rows_ok(C_dense) || !rows_ok(c_val))
        throw py::value_error("leading dimensions must be nsys or 1 (shared)");
    py::array_t<float> lam({nsys, kNl}), dz({nsys, kNz});
    py::array_t<int> iters(nsys);
    auto ld = [&](const farr& a) -> long long { return a.shape(0) == 1 && nsys > 1 ? 0 : (long long)a.shape(1); };
    const long long lG = ld(G_dense), lC = ld(C_dense), lg = ld(g_val), lc = ld(c_val);
    const float *pG = G_dense.data(), *pC = C_dense.data(), *pg = g_val.data(), *pc = c_val.data();
    float* pl_ = lam.mutable_data(); float* pd = dz.mutable_data(); int* pi = iters.mutable_data();
    Err err;
    locked_nogil([&] {
        if (!g_block_plan) {
            int dev = 0;
            if (const char* d = std::getenv("GATO_B200_DEVICE")) dev = std::atoi(d);
            err.set(gato_plan_create_blocks(kS, kC, kK, dev, 1, &g_block_plan));
            if (err.rc) { g_block_plan = nullptr; return; }
        }
        float ms = 0.f;
        err.set(gato_solve_batched(g_block_plan, nsys, pG, lG, pC, lC, pg, lg, pc, lc, nullptr, 0, pl_, pd, pi,
                                   exit_tol, max_iters, GATO_WARM_OFF, rho, &ms));
        g_cache.avg_ms = g_cache.first_ms = ms;
    });
    err.raise();
    return py::make_tuple(lam, dz, iters);
}

// EXT: MPC sequence.  g_val [T, n_z], c_val [T, n_l], rho [T]; G_val / C_val [T or 1, nnz].
// warm_start=True chains lambda on the device (a TRUE warm start, r0 = gamma - S lambda_{t-1}); False
// starts every step cold, which is also what the reference does with warm_start=True.
py::tuple linsys_solve_sequence(iarr G_row, iarr G_col, farr G_val, iarr C_row, iarr C_col, farr C_val,
                                farr g_val, farr c_val, farr rho, float exit_tol, int max_iters, bool warm_start) {
    if (G_val.ndim() != 2 || C_val.ndim() != 2 || g_val.ndim() != 2 || c_val.ndim() != 2 || rho.ndim() != 1)
        throw py::value_error("G_val, C_val, g_val, c_val must be [T, n] arrays and rho [T]");
    const int T = (int)g_val.shape(0);
    check_sizes(G_row.size(), G_col.size(), G_val.shape(1), C_row.size(), C_col.size(), C_val.shape(1),
                g_val.shape(1), c_val.shape(1), kNl);
    auto rows_ok = [&](const farr& a) { return a.shape(0) == T || a.shape(0) == 1; };
    if (!rows_ok(G_val) || !rows_ok(C_val) || !rows_ok(c_val) || rho.shape(0) != T)
        throw py::value_error("leading dimensions must be T or 1 (shared); rho must have T entries");
    std::vector<int> gr(G_row.data(), G_row.data() + G_row.size()), gc(G_col.data(), G_col.data() + G_col.size());
    std::vector<int> cr(C_row.data(), C_row.data() + C_row.size()), cc(C_col.data(), C_col.data() + C_col.size());
    py::array_t<float> lam({T, kNl}), dz({T, kNz});
    py::array_t<int> iters(T);
    auto ld = [&](const farr& a) -> long long { return a.shape(0) == 1 && T > 1 ? 0 : (long long)a.shape(1); };
    const long long lG = ld(G_val), lC = ld(C_val), lg = ld(g_val), lc = ld(c_val);
    const float *pG = G_val.data(), *pC = C_val.data(), *pg = g_val.data(), *pc = c_val.data(), *pr = rho.data();
    float* pl_ = lam.mutable_data(); float* pd = dz.mutable_data(); int* pi = iters.mutable_data();
    Err err;
    locked_nogil([&] {
        gato_plan* plan = plan_for(gr, gc, cr, cc, &err);
        if (!plan) return;
        float ms = 0.f;
        err.set(gato_solve_sequence(plan, T, pG, lG, pC, lC, pg, lg, pc, lc, pr, 0.f, nullptr, pl_, pd, pi, exit_tol,
                                    max_iters, warm_start ? GATO_WARM_TRUE : GATO_WARM_OFF, &ms));
        g_cache.avg_ms = ms / (float)T;
        g_cache.first_ms = ms;
    });
    err.raise();
    return py::make_tuple(lam, dz, iters);
}

}  // namespace

PYBIND11_MODULE(gpu_library, m) {
    m.doc() = "B200-native GATO KKT linear-system solve (drop-in for the reference gpu_library)";
    m.attr("STATE_SIZE") = kS;
    m.attr("CONTROL_SIZE") = kC;
    m.attr("KNOT_POINTS") = kK;
    m.def("linsys_solve", &linsys_solve);   // positional, like the reference (no py::arg names)
    m.def("linsys_solve_info", &linsys_solve_info);
    m.def("linsys_solve_batched", &linsys_solve_batched);
    m.def("linsys_solve_blocks", &linsys_solve_blocks);
    m.def("linsys_solve_sequence", &linsys_solve_sequence);
    m.def("set_true_warm_start", [](bool on) { locked_nogil([&] { g_cache.true_warm = on; }); });
    m.def("set_cluster", [](int cl) {
        Err err;
        locked_nogil([&] { if (g_cache.plan) err.set(gato_plan_set_cluster(g_cache.plan, cl)); });
        err.raise();
    });
}
