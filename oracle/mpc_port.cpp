// TEST INFRASTRUCTURE -- plain C++ (float64, single thread per problem) restatement of the reference's MPC step.
//
// A second, independent CPU path next to oracle/mpc_oracle.py: it follows the reference line by line
//   nearest index        main/lib/trajectories.py:100-126          reference lookup   main/lib/mpc.py:86-109
//   rollout              main/lib/mpc.py:112-126, simulation.py:35-47, bicycle/main.py:28-41
//   linearisation        main/lib/mpc.py:58-79                      cost / constraints main/lib/mpc.py:138-191
//   bookkeeping          main/lib/mpc.py:211-239, 280-299
// and replaces the cvxpy -> ECOS call (mpc.py:193-194; absent, unpinned) by a dense Mehrotra interior point on the
// problem condensed with GENERIC dense algebra (explicit 4x4 / 4x2 model matrices multiplied out, dense G, dense
// Cholesky) -- none of the closed forms, scans or structured operators of the CUDA kernels.  It is what
// bench.py times as `cpu_baseline` / `--impl reference` (kind "port") and what tests/test_cpp_port.py pins against
// the golden vectors recorded from the reference.  Only tests/, smoke() and bench.py's CPU legs may load it.
//
// build: g++ -O2 -shared -fPIC -std=c++17 -pthread oracle/mpc_port.cpp -o oracle/_build/libmpcport.so
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct PortConfig {          // plain mirror of the fields used (filled from Python)
    int T, max_iter;
    double w_perp, w_para, R[2], Rd[2], Q_v_yaw[2], Qf_scaled[4];
    double max_dsteer, max_accel, max_decel, max_steer, max_speed, min_speed, r_end, min_ref_speed, dt;
};

using Vec = std::vector<double>;
struct Mat {
    int r = 0, c = 0;
    Vec a;
    Mat() {}
    Mat(int r_, int c_) : r(r_), c(c_), a((size_t)r_ * c_, 0.0) {}
    double& operator()(int i, int j) { return a[(size_t)i * c + j]; }
    double operator()(int i, int j) const { return a[(size_t)i * c + j]; }
};

// main/lib/trajectories.py:100-126 (forward=True); -1 where the reference raises
int nearest_index(const double* cx, const double* cy, int nc, int start, double x, double y) {
    const int n = nc - start;
    if (n >= 3) {
        std::vector<std::pair<double, int>> d(n);
        for (int i = 0; i < n; ++i) {
            const double dx = cx[start + i] - x, dy = cy[start + i] - y;
            d[i] = {std::sqrt(dx * dx + dy * dy), i};
        }
        std::partial_sort(d.begin(), d.begin() + 3, d.end());          // (distance, index): stable on ties
        const int i0 = d[0].second, i1 = d[1].second, i2 = d[2].second;
        if (std::abs(i1 - i2) == 2) return i0 + start;
        if (std::abs(i0 - i1) == 1) return std::max(i0, i1) + start;
        return -1;
    }
    if (n == 2) return 1 + start;
    return start;
}

// in-place Cholesky A = L L' (lower); false if not positive definite
bool cholesky(Mat& A) {
    const int n = A.r;
    for (int j = 0; j < n; ++j) {
        double d = A(j, j);
        for (int k = 0; k < j; ++k) d -= A(j, k) * A(j, k);
        if (!(d > 0.0)) return false;
        d = std::sqrt(d);
        A(j, j) = d;
        for (int i = j + 1; i < n; ++i) {
            double s = A(i, j);
            for (int k = 0; k < j; ++k) s -= A(i, k) * A(j, k);
            A(i, j) = s / d;
        }
    }
    return true;
}
void chol_solve(const Mat& L, Vec& b) {
    const int n = L.r;
    for (int i = 0; i < n; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s -= L(i, k) * b[k];
        b[i] = s / L(i, i);
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < n; ++k) s -= L(k, i) * b[k];
        b[i] = s / L(i, i);
    }
}

// min 1/2 u'Hu + g'u  s.t.  G u <= h  (H SPD): Mehrotra predictor-corrector, dense normal equations.
// returns 0 optimal, 1 inaccurate, 2 failed
int ipm(const Mat& H, const Vec& g, const Mat& G, const Vec& h, Vec& u, int& iters) {
    const int n = H.r, m = G.r;
    Vec s(m), lam(m, 1.0), rd(n), rp(m), w(m), Gu(m);
    auto mulG = [&](const Vec& x, Vec& y) { for (int i = 0; i < m; ++i) { double t = 0; for (int j = 0; j < n; ++j) t += G(i, j) * x[j]; y[i] = t; } };
    auto mulGT = [&](const Vec& y, Vec& x) { std::fill(x.begin(), x.end(), 0.0); for (int i = 0; i < m; ++i) { const double t = y[i]; if (t != 0.0) for (int j = 0; j < n; ++j) x[j] += G(i, j) * t; } };
    mulG(u, Gu);
    for (int i = 0; i < m; ++i) s[i] = std::max(h[i] - Gu[i], 1.0);
    double gscale = 1.0;
    for (int j = 0; j < n; ++j) gscale = std::max(gscale, 1.0 + std::fabs(g[j]));
    double best_merit = 1e300;
    Vec best_u = u;
    int status = 2;
    Vec tmpn(n), y(m), du(n), ds(m), dl(m), rc(m), rhs(n);
    Mat M(n, n);
    int it = 0;
    for (; it < 80; ++it) {
        mulG(u, Gu);
        mulGT(lam, tmpn);
        double nrd = 0, nrp = 0, mu = 0;
        for (int j = 0; j < n; ++j) { double t = g[j] + tmpn[j]; for (int k = 0; k < n; ++k) t += H(j, k) * u[k]; rd[j] = t; nrd = std::max(nrd, std::fabs(t)); }
        for (int i = 0; i < m; ++i) { rp[i] = Gu[i] + s[i] - h[i]; nrp = std::max(nrp, std::fabs(rp[i])); mu += s[i] * lam[i]; }
        mu /= m;
        const double merit = std::max({nrd / gscale, nrp, mu});
        if (merit < best_merit) { best_merit = merit; best_u = u; }
        if (nrd < 1e-9 + 1e-12 * gscale && nrp < 1e-10 && mu < 1e-12) { status = 0; break; }
        if (!(mu == mu) || mu < 1e-16) break;
        for (int i = 0; i < m; ++i) w[i] = lam[i] / s[i];
        for (int a = 0; a < n; ++a) for (int b = 0; b <= a; ++b) M(a, b) = H(a, b);
        for (int i = 0; i < m; ++i) {
            const double wi = w[i];
            for (int a = 0; a < n; ++a) { const double ga = G(i, a) * wi; if (ga != 0.0) for (int b = 0; b <= a; ++b) M(a, b) += ga * G(i, b); }
        }
        if (!cholesky(M)) break;
        auto direction = [&](const Vec& rcv) {
            for (int i = 0; i < m; ++i) y[i] = rcv[i] / s[i] - w[i] * rp[i];
            mulGT(y, rhs);
            for (int j = 0; j < n; ++j) rhs[j] -= rd[j];
            du = rhs;
            chol_solve(M, du);
            mulG(du, Gu);
            for (int i = 0; i < m; ++i) { ds[i] = -rp[i] - Gu[i]; dl[i] = -(rcv[i] + lam[i] * ds[i]) / s[i]; }
        };
        auto maxstep = [&]() { double t = 0; for (int i = 0; i < m; ++i) t = std::max({t, -ds[i] / s[i], -dl[i] / lam[i]}); return t; };
        for (int i = 0; i < m; ++i) rc[i] = s[i] * lam[i];
        direction(rc);
        const double al_aff = 1.0 / std::max(1.0, maxstep());
        double mu_aff = 0;
        for (int i = 0; i < m; ++i) mu_aff += (s[i] + al_aff * ds[i]) * (lam[i] + al_aff * dl[i]);
        mu_aff /= m;
        const double sig = std::pow(mu_aff / mu, 3.0);
        for (int i = 0; i < m; ++i) rc[i] = s[i] * lam[i] + ds[i] * dl[i] - sig * mu;
        direction(rc);
        const double tm = maxstep();
        const double al = tm > 0.99 ? 0.99 / tm : 1.0;
        for (int j = 0; j < n; ++j) u[j] += al * du[j];
        for (int i = 0; i < m; ++i) { s[i] += al * ds[i]; lam[i] += al * dl[i]; }
    }
    iters = it;
    if (status != 0) { u = best_u; status = best_merit < 1e-6 ? 1 : 2; }
    return status;
}

struct StepIO {
    // in
    const double* cx; const double* cy; const double* cyaw; int nc; double dl, L;
    double x, y, yaw, v;
    int target_ind, warm;
    // io
    double* oa; double* od;              // [T]
    // out
    double* traj;                        // [4][T+1] ox oy ov oyaw (may be null)
    double* xref;                        // [4][T+1] (may be null)
    double di, ai; int status, iters;
};

void mpc_step(const PortConfig& c, StepIO& io) {
    const int T = c.T, n = 2 * T;
    const double dt = c.dt;
    Vec oa(T, 0.0), od(T, 0.0);
    if (io.warm) { for (int t = 0; t < T; ++t) { oa[t] = io.oa[t]; od[t] = io.od[t]; } }
    Vec ov;                               // previous pass's speeds (MAX_ITER > 1)
    bool solved = false;
    Vec sol_u(n), X(4 * (T + 1));
    Vec xr(4 * (T + 1), 0.0);
    io.status = 0; io.iters = 0;
    for (int pass = 0; pass < c.max_iter; ++pass) {
        // ---- _calc_ref_trajectory (mpc.py:86-109)
        const int ni = nearest_index(io.cx, io.cy, io.nc, io.target_ind, io.x, io.y);
        if (ni < 0) { io.status = 3; return; }                   // the reference raises: nothing changes
        io.target_ind = ni;
        std::vector<int> idx(T + 1);
        std::vector<char> reach(T + 1);
        double trav = 0.0;
        for (int t = 0; t <= T; ++t) {
            const double sp = ov.empty() ? std::max(io.v, c.min_ref_speed) : ov[t];
            trav = (t == 0) ? std::fabs(sp) * dt : trav + std::fabs(sp) * dt;      // np.cumsum
            int k = (int)std::nearbyint(trav / io.dl) + ni;                        // np.rint: half to even
            k = std::min(k, io.nc - 1);
            idx[t] = k; reach[t] = (k == io.nc - 1);
            xr[0 * (T + 1) + t] = io.cx[k]; xr[1 * (T + 1) + t] = io.cy[k]; xr[2 * (T + 1) + t] = 0.0; xr[3 * (T + 1) + t] = io.cyaw[k];
        }
        // ---- _predict_motion (mpc.py:112-126): x, y, v, yaw
        Vec xb(4 * (T + 1));
        {
            double x = io.x, y = io.y, v = io.v, yaw = io.yaw;
            xb[0] = x; xb[T + 1] = y; xb[2 * (T + 1)] = v; xb[3 * (T + 1)] = yaw;
            for (int t = 1; t <= T; ++t) {
                const double d = std::max(std::min(od[t - 1], c.max_steer), -c.max_steer);
                const double xd = v * std::cos(yaw), yd = v * std::sin(yaw), td = (v / io.L) * std::tan(d);
                x = x + xd * dt; y = y + yd * dt; yaw = yaw + td * dt;
                v = v + oa[t - 1] * dt; v = std::max(std::min(v, c.max_speed), c.min_speed);
                xb[t] = x; xb[T + 1 + t] = y; xb[2 * (T + 1) + t] = v; xb[3 * (T + 1) + t] = yaw;
            }
        }
        // ---- condensing with explicit model matrices (mpc.py:58-79, 173-175): x_t = Phi_t x0 + Gam_t u + c_t
        std::vector<Mat> Gam(T + 1, Mat(4, n));
        std::vector<Vec> free_(T + 1, Vec(4, 0.0));
        free_[0] = {io.x, io.y, io.v, io.yaw};
        for (int t = 0; t < T; ++t) {
            const double v = xb[2 * (T + 1) + t], phi = xb[3 * (T + 1) + t], delta = 0.0;
            double A[4][4] = {{1, 0, dt * std::cos(phi), -dt * v * std::sin(phi)}, {0, 1, dt * std::sin(phi), dt * v * std::cos(phi)},
                              {0, 0, 1, 0}, {0, 0, dt * std::tan(delta) / io.L, 1}};
            double Bm[4][2] = {{0, 0}, {0, 0}, {dt, 0}, {0, dt * v / (io.L * std::cos(delta) * std::cos(delta))}};
            double Cv[4] = {dt * v * std::sin(phi) * phi, -dt * v * std::cos(phi) * phi, 0.0, -dt * v * delta / (io.L * std::cos(delta) * std::cos(delta))};
            for (int i = 0; i < 4; ++i) {
                double f = Cv[i];
                for (int k = 0; k < 4; ++k) f += A[i][k] * free_[t][k];
                free_[t + 1][i] = f;
                for (int j = 0; j < n; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += A[i][k] * Gam[t](k, j); Gam[t + 1](i, j) = s; }
                Gam[t + 1](i, t) += Bm[i][0];            // u index: a_t = t, delta_t = T + t
                Gam[t + 1](i, T + t) += Bm[i][1];
            }
        }
        // ---- cost (mpc.py:151-184): sum quad_form (no 1/2) -> H = 2 sum Gam' Q Gam + ..., g = 2 sum Gam' Q (free - ref)
        Mat H(n, n); Vec g(n, 0.0);
        for (int t = 1; t <= T; ++t) {
            double Q[4][4] = {{0}};
            if (!reach[t]) {
                const double psi = xr[3 * (T + 1) + t];
                const double c1 = std::cos(psi + 0.5 * M_PI), s1 = std::sin(psi + 0.5 * M_PI), c2 = std::cos(psi), s2 = std::sin(psi);
                Q[0][0] = c.w_perp * c1 * c1 + c.w_para * c2 * c2; Q[0][1] = Q[1][0] = c.w_perp * c1 * s1 + c.w_para * c2 * s2;
                Q[1][1] = c.w_perp * s1 * s1 + c.w_para * s2 * s2; Q[2][2] = c.Q_v_yaw[0]; Q[3][3] = c.Q_v_yaw[1];
            } else { for (int i = 0; i < 4; ++i) Q[i][i] = c.Qf_scaled[i]; }
            double e[4], Qe[4];
            for (int i = 0; i < 4; ++i) e[i] = free_[t][i] - xr[i * (T + 1) + t];
            for (int i = 0; i < 4; ++i) { Qe[i] = 0; for (int k = 0; k < 4; ++k) Qe[i] += Q[i][k] * e[k]; }
            Mat QG(4, n);
            for (int i = 0; i < 4; ++i) for (int j = 0; j < n; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += Q[i][k] * Gam[t](k, j); QG(i, j) = s; }
            for (int a = 0; a < n; ++a) {
                double ga = 0;
                for (int i = 0; i < 4; ++i) ga += Gam[t](i, a) * Qe[i];
                g[a] += 2.0 * ga;
                for (int b = 0; b < n; ++b) { double s = 0; for (int i = 0; i < 4; ++i) s += Gam[t](i, a) * QG(i, b); H(a, b) += 2.0 * s; }
            }
        }
        for (int t = 0; t < T; ++t) {
            const double ra = reach[t] ? c.r_end : c.R[0], rd = reach[t] ? c.r_end : c.R[1];      // mpc.py:177-180
            H(t, t) += 2.0 * ra; H(T + t, T + t) += 2.0 * rd;
            if (t < T - 1) {                                                                     // mpc.py:183
                for (int blk = 0; blk < 2; ++blk) {
                    const int i0 = blk * T + t, i1 = i0 + 1; const double r = c.Rd[blk];
                    H(i0, i0) += 2.0 * r; H(i1, i1) += 2.0 * r; H(i0, i1) -= 2.0 * r; H(i1, i0) -= 2.0 * r;
                }
            }
        }
        // ---- constraints (mpc.py:184,186-191); x[2,0] rows are constants
        if (io.v > c.max_speed + 1e-9 || io.v < c.min_speed - 1e-9) { solved = false; io.status = 2; break; }
        const int m = 8 * T - 2;
        Mat G(m, n); Vec hv(m);
        int r = 0;
        for (int t = 0; t < T; ++t) { G(r, t) = 1; hv[r++] = c.max_accel; G(r, t) = -1; hv[r++] = -c.max_decel; }
        for (int t = 0; t < T; ++t) { G(r, T + t) = 1; hv[r++] = c.max_steer; G(r, T + t) = -1; hv[r++] = c.max_steer; }
        for (int t = 0; t < T - 1; ++t) { G(r, T + t + 1) = 1; G(r, T + t) = -1; hv[r++] = c.max_dsteer * dt; G(r, T + t + 1) = -1; G(r, T + t) = 1; hv[r++] = c.max_dsteer * dt; }
        for (int t = 1; t <= T; ++t) {
            for (int j = 0; j < n; ++j) { G(r, j) = Gam[t](2, j); G(r + 1, j) = -Gam[t](2, j); }
            hv[r] = c.max_speed - free_[t][2]; hv[r + 1] = free_[t][2] - c.min_speed; r += 2;
        }
        Vec u(n);
        for (int t = 0; t < T; ++t) { u[t] = oa[t]; u[T + t] = od[t]; }
        int its = 0;
        const int st = ipm(H, g, G, hv, u, its);
        io.iters = its;
        if (st == 2) { solved = false; io.status = 2; break; }
        io.status = st; solved = true; sol_u = u;
        for (int t = 0; t <= T; ++t) for (int i = 0; i < 4; ++i) { double s = free_[t][i]; for (int j = 0; j < n; ++j) s += Gam[t](i, j) * u[j]; X[i * (T + 1) + t] = s; }
        for (int t = 0; t < T; ++t) { oa[t] = u[t]; od[t] = u[T + t]; }
        ov.assign(X.begin() + 2 * (T + 1), X.begin() + 3 * (T + 1));
    }
    if (io.xref) std::memcpy(io.xref, xr.data(), sizeof(double) * 4 * (T + 1));
    if (solved) {                                   // mpc.py:289-293
        for (int t = 0; t < T; ++t) { io.oa[t] = sol_u[t]; io.od[t] = sol_u[T + t]; }
        if (io.traj) std::memcpy(io.traj, X.data(), sizeof(double) * 4 * (T + 1));
        io.di = sol_u[T]; io.ai = sol_u[0]; io.warm = 1;
    } else {                                        // mpc.py:294-297
        io.ai = c.max_decel; io.warm = 0;
        if (io.traj) for (int i = 0; i < 4 * (T + 1); ++i) io.traj[i] = NAN;
    }
}

}  // namespace

extern "C" {

// cfg: 24 doubles [T, max_iter, w_perp, w_para, R0, R1, Rd0, Rd1, Qv, Qyaw, Qf0..3 (scaled by T), max_dsteer(rad/s),
//      max_accel, max_decel, max_steer, max_speed, min_speed, r_end, min_ref_speed, dt, 0]
// courses [R][3][ncmax]; per problem: state [B,4] (x,y,yaw,v), route, ncourse, L, target_ind (io), warm (io), oa/od [B,T] (io),
// traj/xref [B,4,T+1] (out, may be null), di_ai [B,2] (io), status/iters [B] (out).  nthreads <= 0: hardware concurrency.
int mpcport_step_batch(const double* cfg, const double* courses, const double* dl, int ncmax, int B, const double* state,
                       const int* route, const int* ncourse, const double* L, int* target_ind, int* warm, double* oa, double* od,
                       double* traj, double* xref, double* di_ai, int* status, int* iters, int nthreads) {
    PortConfig c;
    c.T = (int)cfg[0]; c.max_iter = (int)cfg[1]; c.w_perp = cfg[2]; c.w_para = cfg[3]; c.R[0] = cfg[4]; c.R[1] = cfg[5];
    c.Rd[0] = cfg[6]; c.Rd[1] = cfg[7]; c.Q_v_yaw[0] = cfg[8]; c.Q_v_yaw[1] = cfg[9];
    for (int i = 0; i < 4; ++i) c.Qf_scaled[i] = cfg[10 + i];
    c.max_dsteer = cfg[14]; c.max_accel = cfg[15]; c.max_decel = cfg[16]; c.max_steer = cfg[17]; c.max_speed = cfg[18];
    c.min_speed = cfg[19]; c.r_end = cfg[20]; c.min_ref_speed = cfg[21]; c.dt = cfg[22];
    const int T = c.T;
    if (nthreads <= 0) nthreads = (int)std::max(1u, std::thread::hardware_concurrency());
    nthreads = std::min(nthreads, std::max(B, 1));
    auto work = [&](int t0) {
        for (int b = t0; b < B; b += nthreads) {
            StepIO io;
            const double* cr = courses + (size_t)route[b] * 3 * ncmax;
            io.cx = cr; io.cy = cr + ncmax; io.cyaw = cr + 2 * (size_t)ncmax; io.nc = ncourse[b]; io.dl = dl[route[b]]; io.L = L[b];
            io.x = state[4 * b]; io.y = state[4 * b + 1]; io.yaw = state[4 * b + 2]; io.v = state[4 * b + 3];
            io.target_ind = target_ind[b]; io.warm = warm[b];
            io.oa = oa + (size_t)b * T; io.od = od + (size_t)b * T;
            io.traj = traj ? traj + (size_t)b * 4 * (T + 1) : nullptr; io.xref = xref ? xref + (size_t)b * 4 * (T + 1) : nullptr;
            io.di = di_ai[2 * b]; io.ai = di_ai[2 * b + 1];
            mpc_step(c, io);
            status[b] = io.status; iters[b] = io.iters;
            if (io.status == 3) continue;
            target_ind[b] = io.target_ind; warm[b] = io.warm; di_ai[2 * b] = io.di; di_ai[2 * b + 1] = io.ai;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    return 0;
}

int mpcport_version(void) { return 1; }

}  // extern "C"
