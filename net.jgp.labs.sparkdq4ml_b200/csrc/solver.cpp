// solver.cpp -- driver-side optimisers and normal-equation solvers (see solver.hpp for what they follow).
#include "solver.hpp"

#include <algorithm>
#include <cmath>
#include <deque>
#include <limits>

#include "../../include/dq4ml.h"

namespace dq4ml {
namespace {

inline double dot(const Vec& a, const Vec& b) {
  double s = 0.0;
  for (size_t i = 0; i < a.size(); ++i) s += a[i] * b[i];
  return s;
}
inline double norm2(const Vec& a) { return std::sqrt(dot(a, a)); }
inline void axpy(double alpha, const Vec& x, Vec& y) {
  for (size_t i = 0; i < x.size(); ++i) y[i] += alpha * x[i];
}
inline double sgn(double v) { return v > 0.0 ? 1.0 : (v < 0.0 ? -1.0 : 0.0); }

struct QnFailure {
  const char* what;
};

// One-slot memo in front of the objective (Breeze CachedDiffFunction): the key is the vector as it looks AFTER the
// call, because the objective may edit it.
class MemoObjective {
 public:
  explicit MemoObjective(const CostFn& f) : f_(f) {}
  double operator()(Vec& x, Vec& grad) {
    if (!have_ || x != key_) {
      val_ = f_(x, grad_);
      key_ = x;
      have_ = true;
      ++evals_;
    }
    grad = grad_;
    return val_;
  }
  int evals() const { return evals_; }

 private:
  const CostFn& f_;
  bool have_ = false;
  Vec key_, grad_;
  double val_ = 0.0;
  int evals_ = 0;
};

struct Iterate {
  Vec x, grad, adj_grad;
  double value = 0.0, adj_value = 0.0;
  int iter = 0;
  std::deque<Vec> steps, grad_deltas;  // newest first
  std::deque<double> recent_values{std::numeric_limits<double>::infinity()};
  bool search_failed = false;
};

class QuasiNewton {
 public:
  QuasiNewton(const CostFn& f, int max_iter, double tol, const L1Fn& l1, int m)
      : f_(f), max_iter_(max_iter), tol_(tol), l1_(l1), m_(m), owlqn_(static_cast<bool>(l1)) {}

  QnResult run(Vec x0) {
    QnResult res;
    Iterate s;
    s.x = std::move(x0);
    s.value = f_(s.x, s.grad);
    s.adj_value = regularise(s.x, s.grad, s.value, s.adj_grad);
    const double initial_adj = s.adj_value;
    bool failed_once = false;
    for (;;) {
      res.adjusted_values.push_back(s.adj_value);
      if (const char* why = stop_reason(s, initial_adj)) {
        res.stop_reason = why;
        break;
      }
      try {
        Vec dir = descent_direction(s);
        const double alpha = owlqn_ ? backtracking_step(s, dir) : strong_wolfe_step(s, dir);
        Iterate nx;
        nx.x = step_from(s, dir, alpha);
        nx.value = f_(nx.x, nx.grad);
        nx.adj_value = regularise(nx.x, nx.grad, nx.value, nx.adj_grad);
        nx.iter = s.iter + 1;
        nx.steps = s.steps;
        nx.grad_deltas = s.grad_deltas;
        Vec sk(nx.x.size()), yk(nx.x.size());
        for (size_t i = 0; i < sk.size(); ++i) {
          sk[i] = nx.x[i] - s.x[i];
          yk[i] = nx.grad[i] - s.grad[i];
        }
        nx.steps.push_front(std::move(sk));
        nx.grad_deltas.push_front(std::move(yk));
        if (static_cast<int>(nx.steps.size()) > m_) {
          nx.steps.pop_back();
          nx.grad_deltas.pop_back();
        }
        nx.recent_values = s.recent_values;
        nx.recent_values.push_back(nx.value);
        while (nx.recent_values.size() > kValueMemory) nx.recent_values.pop_front();
        failed_once = false;
        s = std::move(nx);
      } catch (const QnFailure&) {
        if (!failed_once) {  // first failure: forget the curvature pairs and retry from the same point
          failed_once = true;
          s.steps.clear();
          s.grad_deltas.clear();
        } else {
          s.search_failed = true;
        }
      }
    }
    res.x = s.x;
    res.evaluations = f_.evals();
    return res;
  }

 private:
  static constexpr size_t kValueMemory = 20;  // FirstOrderMinimizer.defaultConvergenceCheck fvalMemory

  const char* stop_reason(const Iterate& s, double initial_adj) const {
    if (max_iter_ >= 0 && s.iter >= max_iter_) return "max iterations";
    if (s.recent_values.size() >= 2) {
      const double mx = *std::max_element(s.recent_values.begin(), s.recent_values.end());
      if (std::fabs(s.adj_value - mx) <= tol_ * initial_adj) return "function values converged";
    }
    if (norm2(s.adj_grad) <= std::max(tol_ * s.adj_value, 1e-8)) return "gradient converged";
    if (s.search_failed) return "line search failed";
    return nullptr;
  }

  // OWL-QN: pseudo-gradient and value + sum |lambda_i x_i|.  Plain L-BFGS: identity.
  double regularise(const Vec& x, const Vec& grad, double value, Vec& adj_grad) const {
    adj_grad = grad;
    if (!owlqn_) return value;
    double adj = value;
    for (size_t i = 0; i < x.size(); ++i) {
      const double lam = l1_(static_cast<int>(i));
      if (lam == 0.0) continue;
      adj += std::fabs(lam * x[i]);
      if (x[i] == 0.0) {
        const double up = grad[i] + lam, dn = grad[i] - lam;
        adj_grad[i] = dn > 0.0 ? dn : (up < 0.0 ? up : 0.0);
      } else {
        adj_grad[i] = grad[i] + sgn(x[i]) * lam;
      }
    }
    return adj;
  }

  // Two-loop recursion with H0 = (s.y)/(y.y) of the newest pair, applied to the (pseudo-)gradient and negated;
  // OWL-QN then zeroes components that do not oppose the pseudo-gradient.
  Vec descent_direction(const Iterate& s) const {
    const Vec& g = owlqn_ ? s.adj_grad : s.grad;
    const size_t h = s.steps.size();
    double h0 = 1.0;
    if (h > 0) {
      const double sy = dot(s.steps[0], s.grad_deltas[0]);
      const double yy = dot(s.grad_deltas[0], s.grad_deltas[0]);
      if (sy < 0.0 || std::isnan(sy)) throw QnFailure{"NaN history"};
      h0 = sy / yy;
    }
    Vec dir = g;
    Vec a(h), rho(h);
    for (size_t i = 0; i < h; ++i) {
      rho[i] = dot(s.steps[i], s.grad_deltas[i]);
      a[i] = dot(s.steps[i], dir) / rho[i];
      if (std::isnan(a[i])) throw QnFailure{"NaN history"};
      axpy(-a[i], s.grad_deltas[i], dir);
    }
    for (double& v : dir) v *= h0;
    for (size_t k = h; k-- > 0;) {
      const double beta = dot(s.grad_deltas[k], dir) / rho[k];
      axpy(a[k] - beta, s.steps[k], dir);
    }
    for (double& v : dir) v *= -1.0;
    if (owlqn_)
      for (size_t i = 0; i < dir.size(); ++i)
        if (!(dir[i] * s.adj_grad[i] < 0.0)) dir[i] = 0.0;
    return dir;
  }

  // x + alpha*dir; OWL-QN additionally clamps coordinates that leave the orthant chosen at s to zero.
  Vec step_from(const Iterate& s, const Vec& dir, double alpha) const {
    Vec nx(s.x.size());
    for (size_t i = 0; i < nx.size(); ++i) nx[i] = s.x[i] + dir[i] * alpha;
    if (owlqn_) {
      for (size_t i = 0; i < nx.size(); ++i) {
        const double orth = s.x[i] != 0.0 ? sgn(s.x[i]) : sgn(-s.adj_grad[i]);
        nx[i] = nx[i] * (sgn(nx[i]) == orth ? 1.0 : 0.0);
      }
    }
    return nx;
  }

  // Breeze BacktrackingLineSearch as OWLQN drives it.  The first trial carries s.value as its function value and
  // only its derivative comes from an evaluation -- a quirk of the library that shapes objectiveHistory[1].
  double backtracking_step(const Iterate& s, const Vec& dir) {
    const double shrink = s.iter < 1 ? 0.1 : 0.5, grow = 2.1, c_armijo = 1e-4, c_wolfe = 0.9;
    const int max_trials = 20;
    auto phi = [&](double alpha, double& deriv) {
      Vec nx = step_from(s, dir, alpha), g, pg;
      const double v = f_(nx, g);
      const double av = regularise(nx, g, v, pg);
      deriv = dot(pg, dir);
      return av;
    };
    double df0, d_init;
    const double f0 = phi(0.0, df0);
    double alpha = s.iter < 1 ? 0.5 / norm2(s.grad) : 1.0;
    phi(alpha, d_init);
    double fval = s.value, fderiv = d_init;
    for (int trial = 0;;) {
      double mult = 1.0;
      if (fval > f0 + alpha * df0 * c_armijo) mult = shrink;
      else if (fderiv < c_wolfe * df0) mult = grow;
      else if (fderiv > -c_wolfe * df0) mult = shrink;
      if (mult == 1.0) return alpha;
      const double next = alpha * mult;
      if (next < 1e-10) throw QnFailure{"step size underflow"};
      if (next > 1e10) throw QnFailure{"step size overflow"};
      const double prev = alpha;
      fval = phi(next, fderiv);
      alpha = next;
      if (++trial >= max_trials) return prev;  // the 21st state is evaluated but never emitted by the iterator
    }
  }

  struct Bracket {
    double t, dd, fval;
  };
  static double cubic(const Bracket& l, const Bracket& r) {
    const double d1 = l.dd + r.dd - 3.0 * (l.fval - r.fval) / (l.t - r.t);
    const double d2 = std::sqrt(d1 * d1 - l.dd * r.dd);
    const double t = r.t - (r.t - l.t) * (r.dd + d2 - d1) / (r.dd - l.dd + 2.0 * d2);
    const double lo = l.t + 0.1 * (r.t - l.t), hi = l.t + 0.9 * (r.t - l.t);
    if (t < lo) return lo;
    if (t > hi) return hi;
    return t;
  }

  // Breeze StrongWolfeLineSearch(maxZoomIter = 10, maxLineSearchIter = 10) as LBFGS drives it.
  double strong_wolfe_step(const Iterate& s, const Vec& dir) {
    const double c1 = 1e-4, c2 = 0.9;
    auto phi = [&](double t) {
      Vec nx(s.x.size()), g;
      for (size_t i = 0; i < nx.size(); ++i) nx[i] = s.x[i] + dir[i] * t;
      Bracket b;
      b.t = t;
      b.fval = f_(nx, g);
      b.dd = dot(g, dir);
      return b;
    };
    double t = s.iter == 0 ? 1.0 / norm2(dir) : 1.0;
    Bracket low = phi(0.0);
    const double fval = low.fval, dd = low.dd;
    if (dd > 0.0) throw QnFailure{"non-descent direction"};
    auto zoom = [&](Bracket lo, Bracket hi) {
      for (int i = 0; i < 10; ++i) {
        const double tt = lo.t > hi.t ? cubic(hi, lo) : cubic(lo, hi);
        const Bracket c = phi(tt);
        if (c.fval > fval + c1 * c.t * dd || c.fval >= lo.fval) {
          hi = c;
        } else {
          if (std::fabs(c.dd) <= c2 * std::fabs(dd)) return c.t;
          if (c.dd * (hi.t - lo.t) >= 0.0) hi = lo;
          lo = c;
        }
      }
      throw QnFailure{"zoom failed"};
    };
    double alpha = -1.0;
    for (int i = 0; i < 10 && alpha < 0.0; ++i) {
      const Bracket c = phi(t);
      if (std::isinf(c.fval) || std::isnan(c.fval)) {
        t /= 2.0;
        continue;
      }
      if (c.fval > fval + c1 * t * dd || (c.fval >= low.fval && i > 0)) alpha = zoom(low, c);
      else if (std::fabs(c.dd) <= c2 * std::fabs(dd)) alpha = c.t;
      else if (c.dd >= 0.0) alpha = zoom(c, low);
      else {
        low = c;
        t *= 1.5;
      }
    }
    if (alpha < 0.0) throw QnFailure{"line search failed"};
    if (alpha * norm2(s.grad) < 1e-10) throw QnFailure{"step size underflow"};
    return alpha;
  }

  MemoObjective f_;
  int max_iter_;
  double tol_;
  const L1Fn& l1_;
  int m_;
  bool owlqn_;
};

// ---------------------------------------------------------------------------------------------------------------------
// Bound-constrained quasi-Newton: breeze.optimize.LBFGSB (Byrd, Lu, Nocedal & Zhu), which LinearRegression.train runs
// for loss = huber so that the scale sigma stays positive.  Hessian approximation in compact form,
// B = theta I - W M W^T with W = [Y, theta S]; every step: generalized Cauchy point along the projected steepest-descent
// path, then (from the second iteration on) an unconstrained minimisation over the variables that are free at the Cauchy
// point, then a strong-Wolfe line search capped at the largest feasible step.  Small dense algebra only (2m <= 20).
struct Dense {  // row-major
  int rows = 0, cols = 0;
  Vec a;
  Dense() = default;
  Dense(int r, int c) : rows(r), cols(c), a(static_cast<size_t>(r) * c, 0.0) {}
  double& at(int i, int j) { return a[static_cast<size_t>(i) * cols + j]; }
  double at(int i, int j) const { return a[static_cast<size_t>(i) * cols + j]; }
};

Vec mat_vec(const Dense& m, const Vec& v) {
  Vec out(m.rows, 0.0);
  for (int i = 0; i < m.rows; ++i) {
    double acc = 0.0;
    for (int j = 0; j < m.cols; ++j) acc += m.at(i, j) * v[j];
    out[i] = acc;
  }
  return out;
}
Vec mat_t_vec(const Dense& m, const Vec& v) {
  Vec out(m.cols, 0.0);
  for (int i = 0; i < m.rows; ++i)
    for (int j = 0; j < m.cols; ++j) out[j] += m.at(i, j) * v[i];
  return out;
}

// Solves A X = B in place by Gaussian elimination with partial pivoting (B holds nrhs columns); false if singular.
bool lu_solve(Dense a, Dense& b) {
  const int n = a.rows;
  for (int k = 0; k < n; ++k) {
    int piv = k;
    for (int i = k + 1; i < n; ++i)
      if (std::fabs(a.at(i, k)) > std::fabs(a.at(piv, k))) piv = i;
    if (a.at(piv, k) == 0.0 || std::isnan(a.at(piv, k))) return false;
    if (piv != k) {
      for (int j = 0; j < n; ++j) std::swap(a.at(k, j), a.at(piv, j));
      for (int j = 0; j < b.cols; ++j) std::swap(b.at(k, j), b.at(piv, j));
    }
    for (int i = k + 1; i < n; ++i) {
      const double f = a.at(i, k) / a.at(k, k);
      if (f == 0.0) continue;
      for (int j = k; j < n; ++j) a.at(i, j) -= f * a.at(k, j);
      for (int j = 0; j < b.cols; ++j) b.at(i, j) -= f * b.at(k, j);
    }
  }
  for (int j = 0; j < b.cols; ++j)
    for (int i = n - 1; i >= 0; --i) {
      double acc = b.at(i, j);
      for (int k = i + 1; k < n; ++k) acc -= a.at(i, k) * b.at(k, j);
      b.at(i, j) = acc / a.at(i, i);
    }
  return true;
}

class BoundedQuasiNewton {
 public:
  BoundedQuasiNewton(const CostFn& f, Vec lower, Vec upper, int max_iter, double tol, int m)
      : f_(f), lo_(std::move(lower)), hi_(std::move(upper)), max_iter_(max_iter), tol_(tol), m_(m) {}

  QnResult run(Vec x0) {
    const int n = static_cast<int>(x0.size());
    QnResult res;
    Point s;
    s.x = std::move(x0);
    s.value = f_(s.x, s.grad);
    reset_curvature(n);
    const double initial = s.value;
    std::deque<double> recent{std::numeric_limits<double>::infinity()};
    bool failed_once = false, search_failed = false;
    for (;;) {
      res.adjusted_values.push_back(s.value);
      if (const char* why = stop_reason(s, recent, initial, search_failed)) {
        res.stop_reason = why;
        break;
      }
      try {
        Vec c;
        Vec target = cauchy_point(s, &c);
        if (s.iter > 0) target = subspace_minimum(s, target, c);
        Vec dir(n);
        for (int i = 0; i < n; ++i) dir[i] = target[i] - s.x[i];
        const double alpha = capped_wolfe_step(s, dir);
        Point nx;
        nx.x = clamped_step(s.x, dir, alpha);
        nx.value = f_(nx.x, nx.grad);
        nx.iter = s.iter + 1;
        Vec sk(n), yk(n);
        for (int i = 0; i < n; ++i) {
          sk[i] = nx.x[i] - s.x[i];
          yk[i] = nx.grad[i] - s.grad[i];
        }
        absorb_pair(sk, yk);
        recent.push_back(nx.value);
        while (recent.size() > 20) recent.pop_front();
        failed_once = false;
        s = std::move(nx);
      } catch (const QnFailure&) {
        if (!failed_once) {
          failed_once = true;
          reset_curvature(n);
        } else {
          search_failed = true;
        }
      }
    }
    res.x = s.x;
    res.evaluations = f_.evals();
    return res;
  }

 private:
  struct Point {
    Vec x, grad;
    double value = 0.0;
    int iter = 0;
  };

  void reset_curvature(int n) {
    theta_ = 1.0;
    w_ = Dense(n, 1);
    mm_ = Dense(1, 1);
    ys_.clear();
    ss_.clear();
  }

  // LBFGSB.defaultConvergenceCheck: projected-gradient step below 1e-5 (infinity norm), or the FirstOrderMinimizer tests
  const char* stop_reason(const Point& s, const std::deque<double>& recent, double initial, bool search_failed) const {
    double pg = 0.0;
    for (size_t i = 0; i < s.x.size(); ++i) {
      double v = s.x[i] - s.grad[i];
      v = v <= lo_[i] ? lo_[i] : (hi_[i] <= v ? hi_[i] : v);
      pg = std::max(pg, std::fabs(v - s.x[i]));
    }
    if (pg < 1e-5) return "projected gradient converged";
    if (max_iter_ >= 0 && s.iter >= max_iter_) return "max iterations";
    if (recent.size() >= 2) {
      const double mx = *std::max_element(recent.begin(), recent.end());
      if (std::fabs(s.value - mx) <= tol_ * initial) return "function values converged";
    }
    if (norm2(s.grad) <= std::max(tol_ * s.value, 1e-8)) return "gradient converged";
    if (search_failed) return "line search failed";
    return nullptr;
  }

  // Algorithm CP.  Returns the Cauchy point; *c_out = W^T (x_cauchy - x).
  Vec cauchy_point(const Point& s, Vec* c_out) const {
    const int n = static_cast<int>(s.x.size());
    const Vec &x = s.x, &g = s.grad;
    Vec d(n, 0.0), t(n, std::numeric_limits<double>::max());
    for (int i = 0; i < n; ++i) {
      if (g[i] == 0.0) continue;
      const double ti = g[i] < 0.0 ? (x[i] - hi_[i]) / g[i] : (x[i] - lo_[i]) / g[i];
      d[i] = ti == 0.0 ? 0.0 : -g[i];
      t[i] = ti;
    }
    Vec xc = x;
    Vec p = mat_t_vec(w_, d);
    Vec c(mm_.rows, 0.0);
    double f1 = dot(g, d);
    double f2 = -theta_ * f1 - dot(p, mat_vec(mm_, p));
    double dt_min = -(f1 / f2);
    double old_t = 0.0;
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return t[a] < t[b]; });
    int i = 0;
    while (i < n && t[order[i]] == 0.0) ++i;
    if (i < n) {
      int b = order[i];
      double min_t = t[b], delta_t = min_t - old_t;
      while (delta_t <= dt_min && i < n) {
        xc[b] = d[b] > 0.0 ? hi_[b] : lo_[b];
        const double zb = xc[b] - x[b];
        for (size_t k = 0; k < c.size(); ++k) c[k] += p[k] * delta_t;
        Vec wb(w_.cols);
        for (int k = 0; k < w_.cols; ++k) wb[k] = w_.at(b, k);
        const double gb = g[b];
        f1 += delta_t * f2 + gb * gb + theta_ * gb * zb - gb * dot(wb, mat_vec(mm_, c));
        f2 += -theta_ * gb * gb - 2.0 * (gb * dot(wb, mat_vec(mm_, p))) - gb * gb * dot(wb, mat_vec(mm_, wb));
        for (size_t k = 0; k < p.size(); ++k) p[k] += wb[k] * gb;
        d[b] = 0.0;
        dt_min = -f1 / f2;
        old_t = min_t;
        ++i;
        if (i < n) {
          b = order[i];
          min_t = t[b];
          delta_t = min_t - old_t;
        }
      }
    }
    dt_min = std::max(dt_min, 0.0);
    old_t += dt_min;
    for (int k = i; k < n; ++k) xc[order[k]] = x[order[k]] + old_t * d[order[k]];
    for (size_t k = 0; k < c.size(); ++k) c[k] += p[k] * dt_min;
    *c_out = c;
    return xc;
  }

  // Direct primal method over the variables that are free at the Cauchy point, shortened by Byrd et al.'s alpha* so that
  // the result stays feasible.  (Breeze 0.13.2's findAlpha, as far as it can be recalled without the jar, never
  // shortens the step; an infeasible subspace minimum then trips `require(init <= bound)` in its line search and the fit
  // dies.  Wherever Breeze succeeds, alpha* = 1 and this is the same algorithm.)
  Vec subspace_minimum(const Point& s, const Vec& xc, const Vec& c) const {
    const int n = static_cast<int>(s.x.size());
    const double inv_theta = 1.0 / theta_;
    std::vector<int> free_idx;
    for (int i = 0; i < n; ++i)
      if (xc[i] != hi_[i] && xc[i] != lo_[i]) free_idx.push_back(i);
    const int nf = static_cast<int>(free_idx.size()), k2 = w_.cols;
    Dense wz(k2, nf);
    for (int j = 0; j < nf; ++j)
      for (int k = 0; k < k2; ++k) wz.at(k, j) = w_.at(free_idx[j], k);
    const Vec mc = mat_vec(mm_, c);
    const Vec wmc = mat_vec(w_, mc);
    Vec rc(nf);
    for (int j = 0; j < nf; ++j) {
      const int i = free_idx[j];
      rc[j] = s.grad[i] + (xc[i] - s.x[i]) * theta_ - wmc[i];
    }
    Vec v = mat_vec(mm_, mat_vec(wz, rc));
    Dense nmat(k2, k2);  // I - M (WZ WZ^T) / theta
    Dense wzwz(k2, k2);
    for (int a = 0; a < k2; ++a)
      for (int b = 0; b < k2; ++b) {
        double acc = 0.0;
        for (int j = 0; j < nf; ++j) acc += wz.at(a, j) * wz.at(b, j);
        wzwz.at(a, b) = acc * inv_theta;
      }
    for (int a = 0; a < k2; ++a)
      for (int b = 0; b < k2; ++b) {
        double acc = 0.0;
        for (int k = 0; k < k2; ++k) acc += mm_.at(a, k) * wzwz.at(k, b);
        nmat.at(a, b) = (a == b ? 1.0 : 0.0) - acc;
      }
    Dense rhs(k2, 1);
    for (int a = 0; a < k2; ++a) rhs.at(a, 0) = v[a];
    if (!lu_solve(nmat, rhs)) throw QnFailure{"singular subspace system"};
    for (int a = 0; a < k2; ++a) v[a] = rhs.at(a, 0);
    const Vec wzv = mat_t_vec(wz, v);
    Vec du(nf);
    for (int j = 0; j < nf; ++j) du[j] = -(rc[j] * inv_theta + wzv[j] * inv_theta * inv_theta);
    double alpha = 1.0;  // the largest step <= 1 along du that stays inside the bounds
    for (int j = 0; j < nf; ++j) {
      const int i = free_idx[j];
      if (du[j] > 0.0) alpha = std::min(alpha, (hi_[i] - xc[i]) / du[j]);
      else if (du[j] < 0.0) alpha = std::min(alpha, (lo_[i] - xc[i]) / du[j]);
    }
    Vec out = xc;
    for (int j = 0; j < nf; ++j) out[free_idx[j]] += alpha * du[j];
    return out;
  }

  void absorb_pair(const Vec& sk, const Vec& yk) {
    std::deque<Vec> ys = ys_, ss = ss_;
    if (static_cast<int>(ys.size()) >= m_) {
      ys.pop_front();
      ss.pop_front();
    }
    ys.push_back(yk);
    ss.push_back(sk);
    const double sy = dot(sk, yk), yy = dot(yk, yk);
    if (!(2.2e-16 * yy < std::fabs(sy))) return;  // curvature too small: the pair is not absorbed at all
    ys_ = std::move(ys);
    ss_ = std::move(ss);
    const int k = static_cast<int>(ys_.size()), n = static_cast<int>(sk.size());
    theta_ = yy / sy;
    w_ = Dense(n, 2 * k);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < k; ++j) {
        w_.at(i, j) = ys_[j][i];
        w_.at(i, k + j) = theta_ * ss_[j][i];
      }
    // M = [[-D, L^T], [L, theta S^T S]]^-1 with A = S^T Y, D = diag(A), L = strictly lower part of A
    Dense big(2 * k, 2 * k);
    for (int a = 0; a < k; ++a)
      for (int b = 0; b < k; ++b) {
        const double sayb = dot(ss_[a], ys_[b]);
        if (a == b) big.at(a, a) = -sayb;
        if (a > b) {
          big.at(k + a, b) = sayb;
          big.at(b, k + a) = sayb;
        }
        big.at(k + a, k + b) = theta_ * dot(ss_[a], ss_[b]);
      }
    Dense inv(2 * k, 2 * k);
    for (int a = 0; a < 2 * k; ++a) inv.at(a, a) = 1.0;
    if (!lu_solve(big, inv)) throw QnFailure{"singular curvature matrix"};
    mm_ = std::move(inv);
  }

  Vec clamped_step(const Vec& x, const Vec& dir, double alpha) const {
    Vec nx(x.size());
    for (size_t i = 0; i < nx.size(); ++i) {
      double v = x[i] + dir[i] * alpha;
      if (v > hi_[i]) v = hi_[i];
      if (v < lo_[i]) v = lo_[i];
      nx[i] = v;
    }
    return nx;
  }

  // StrongWolfeLineSearch(maxZoomIter = 64, maxLineSearchIter = 64).minimizeWithBound(f, 1.0, largest feasible step)
  double capped_wolfe_step(const Point& s, const Vec& dir) {
    double bound = std::numeric_limits<double>::infinity();
    for (size_t i = 0; i < dir.size(); ++i) {
      if (dir[i] == 0.0) continue;
      const double b = ((dir[i] < 0.0 ? lo_[i] : hi_[i]) - s.x[i]) / dir[i];
      if (b < bound) bound = b;
    }
    if (!(1.0 <= bound)) throw SolveError{DQ4ML_ERR_INVALID, "requirement failed: init value should <= bound (LBFGSB line search)"};
    struct Bracket {
      double t, dd, fval;
    };
    const double c1 = 1e-4, c2 = 0.9;
    auto phi = [&](double t) {
      Vec nx = clamped_step(s.x, dir, t), g;
      Bracket b;
      b.t = t;
      b.fval = f_(nx, g);
      b.dd = dot(g, dir);
      return b;
    };
    auto cubic = [](const Bracket& l, const Bracket& r) {
      const double d1 = l.dd + r.dd - 3.0 * (l.fval - r.fval) / (l.t - r.t);
      const double d2 = std::sqrt(d1 * d1 - l.dd * r.dd);
      const double t = r.t - (r.t - l.t) * (r.dd + d2 - d1) / (r.dd - l.dd + 2.0 * d2);
      const double lo = l.t + 0.1 * (r.t - l.t), hi = l.t + 0.9 * (r.t - l.t);
      return t < lo ? lo : (t > hi ? hi : t);
    };
    double t = 1.0;
    Bracket low = phi(0.0);
    const double fval = low.fval, dd = low.dd;
    if (dd > 0.0) throw QnFailure{"non-descent direction"};
    auto zoom = [&](Bracket lo, Bracket hi) {
      for (int i = 0; i < 64; ++i) {
        const double tt = lo.t > hi.t ? cubic(hi, lo) : cubic(lo, hi);
        const Bracket c = phi(tt);
        if (c.fval > fval + c1 * c.t * dd || c.fval >= lo.fval) {
          hi = c;
        } else {
          if (std::fabs(c.dd) <= c2 * std::fabs(dd)) return c.t;
          if (c.dd * (hi.t - lo.t) >= 0.0) hi = lo;
          lo = c;
        }
      }
      throw QnFailure{"zoom failed"};
    };
    for (int i = 0; i < 64; ++i) {
      const Bracket c = phi(t);
      if (std::isinf(c.fval) || std::isnan(c.fval)) {
        t /= 2.0;
        continue;
      }
      if (c.fval > fval + c1 * t * dd || (c.fval >= low.fval && i > 0)) return zoom(low, c);
      if (std::fabs(c.dd) <= c2 * std::fabs(dd)) return c.t;
      if (c.dd >= 0.0) return zoom(c, low);
      low = c;
      if (t == bound) return bound;
      t = std::min(t * 1.5, bound);
    }
    throw QnFailure{"line search failed"};
  }

  MemoObjective f_;
  Vec lo_, hi_;
  int max_iter_;
  double tol_;
  int m_;
  double theta_ = 1.0;
  Dense w_, mm_;
  std::deque<Vec> ys_, ss_;  // oldest first
};

}  // namespace

QnResult lbfgsb_minimize(const CostFn& f, Vec x0, const Vec& lower, const Vec& upper, int max_iter, double tol, int memory) {
  BoundedQuasiNewton qn(f, lower, upper, max_iter, tol, memory);
  return qn.run(std::move(x0));
}

QnResult quasi_newton_minimize(const CostFn& f, Vec x0, int max_iter, double tol, const L1Fn& l1, int memory) {
  QuasiNewton qn(f, max_iter, tol, l1, memory);
  return qn.run(std::move(x0));
}

int packed_cholesky_solve(Vec& ap, Vec& b) {
  const int k = static_cast<int>(b.size());
  auto at = [&](int i, int j) -> double& { return ap[static_cast<size_t>(i) + static_cast<size_t>(j) * (j + 1) / 2]; };
  // factor A = U^T U, column by column
  for (int j = 0; j < k; ++j) {
    double diag = at(j, j);
    for (int i = 0; i < j; ++i) {
      double t = at(i, j);
      for (int l = 0; l < i; ++l) t -= at(l, i) * at(l, j);
      t /= at(i, i);
      at(i, j) = t;
      diag -= t * t;
    }
    if (!(diag > 0.0)) return j + 1;
    at(j, j) = std::sqrt(diag);
  }
  for (int i = 0; i < k; ++i) {  // U^T z = b
    double t = b[i];
    for (int l = 0; l < i; ++l) t -= at(l, i) * b[l];
    b[i] = t / at(i, i);
  }
  for (int i = k - 1; i >= 0; --i) {  // U x = z
    double t = b[i];
    for (int l = i + 1; l < k; ++l) t -= at(i, l) * b[l];
    b[i] = t / at(i, i);
  }
  return 0;
}

bool packed_spd_inverse_diag(Vec ap, int k, Vec* diag) {
  auto at = [&](int i, int j) -> double& { return ap[static_cast<size_t>(i) + static_cast<size_t>(j) * (j + 1) / 2]; };
  // A = U^T U (as in packed_cholesky_solve), then V = U^-1 (upper triangular) and (A^-1)_ii = sum_{j >= i} V_ij^2
  for (int j = 0; j < k; ++j) {
    double d = at(j, j);
    for (int i = 0; i < j; ++i) {
      double t = at(i, j);
      for (int l = 0; l < i; ++l) t -= at(l, i) * at(l, j);
      t /= at(i, i);
      at(i, j) = t;
      d -= t * t;
    }
    if (!(d > 0.0)) return false;
    at(j, j) = std::sqrt(d);
  }
  diag->assign(k, 0.0);
  Vec col(k);
  for (int j = 0; j < k; ++j) {  // column j of V solves U v = e_j by back substitution; only rows <= j are non-zero
    for (int i = j; i >= 0; --i) {
      double t = i == j ? 1.0 : 0.0;
      for (int l = i + 1; l <= j; ++l) t -= at(i, l) * col[l];
      col[i] = t / at(i, i);
    }
    for (int i = 0; i <= j; ++i) (*diag)[i] += col[i] * col[i];
  }
  return true;
}

namespace {

// Regularised incomplete beta function I_x(a, b) by Lentz's continued fraction (Numerical Recipes 6.4).
double beta_cf(double a, double b, double x) {
  const double tiny = 1e-300, eps = 1e-16;
  const double qab = a + b, qap = a + 1.0, qam = a - 1.0;
  double c = 1.0, d = 1.0 - qab * x / qap;
  if (std::fabs(d) < tiny) d = tiny;
  d = 1.0 / d;
  double h = d;
  for (int m = 1; m <= 10000; ++m) {
    const int m2 = 2 * m;
    double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
    d = 1.0 + aa * d;
    if (std::fabs(d) < tiny) d = tiny;
    c = 1.0 + aa / c;
    if (std::fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    h *= d * c;
    aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
    d = 1.0 + aa * d;
    if (std::fabs(d) < tiny) d = tiny;
    c = 1.0 + aa / c;
    if (std::fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    const double del = d * c;
    h *= del;
    if (std::fabs(del - 1.0) < eps) break;
  }
  return h;
}

double reg_inc_beta(double a, double b, double x) {
  if (x <= 0.0) return 0.0;
  if (x >= 1.0) return 1.0;
  const double ln_front = std::lgamma(a + b) - std::lgamma(a) - std::lgamma(b) + a * std::log(x) + b * std::log1p(-x);
  if (x < (a + 1.0) / (a + b + 2.0)) return std::exp(ln_front) * beta_cf(a, b, x) / a;
  return 1.0 - std::exp(ln_front) * beta_cf(b, a, 1.0 - x) / b;
}

}  // namespace

double student_t_two_sided_p(double t, double dof) {
  if (std::isnan(t) || !(dof > 0.0)) return std::numeric_limits<double>::quiet_NaN();
  if (std::isinf(t)) return 0.0;
  return reg_inc_beta(0.5 * dof, 0.5, dof / (dof + t * t));  // = 2 (1 - cdf(|t|))
}

namespace {

// y = A x for packed upper column-major symmetric A of order k.
void packed_symv(int k, const Vec& ap, const Vec& x, Vec& y) {
  y.assign(k, 0.0);
  size_t p = 0;
  for (int j = 0; j < k; ++j) {
    for (int i = 0; i < j; ++i, ++p) {
      y[i] += ap[p] * x[j];
      y[j] += ap[p] * x[i];
    }
    y[j] += ap[p++] * x[j];
  }
}

}  // namespace

LrModel wls_fit_from_moments(const double* mom, int d, const LrParams& p) {
  if (d < 0) throw SolveError{DQ4ML_ERR_INVALID, "negative feature count"};
  const double count = mom[0], w_sum = mom[1];
  const double b_sum = mom[3], bb_sum = mom[4];
  const double* a_sum = mom + 5;
  const double* ab_sum = a_sum + d;
  const double* aa_sum = ab_sum + d;
  const size_t tri = static_cast<size_t>(d) * (d + 1) / 2;
  if (!(count > 0.0) || !(w_sum > 0.0)) throw SolveError{DQ4ML_ERR_EMPTY, "Training dataset is empty."};

  LrModel model;
  model.num_rows = static_cast<int64_t>(count);
  model.data_passes = 1;
  const double raw_b_bar = b_sum / w_sum;
  const double raw_b_std = std::sqrt(std::max(bb_sum / w_sum - raw_b_bar * raw_b_bar, 0.0));
  if (raw_b_std == 0.0) {
    if (p.fit_intercept || raw_b_bar == 0.0) {  // constant label: nothing to train
      model.coefficients.assign(d, 0.0);
      model.intercept = raw_b_bar;
      model.objective_history = {0.0};
      model.solver_used = "trivial";
      return model;
    }
    if (p.reg_param > 0.0)
      throw SolveError{DQ4ML_ERR_INVALID,
                       "The standard deviation of the label is zero. Model cannot be regularized with standardization=true"};
  }
  const double b_std = raw_b_std == 0.0 ? std::fabs(raw_b_bar) : raw_b_std;
  const double b_bar = raw_b_bar / b_std;
  const double bb_bar = (bb_sum / w_sum) / (b_std * b_std);

  Vec a_std(d);
  for (int j = 0; j < d; ++j) {
    const double aw = a_sum[j] / w_sum;
    const double ajj = aa_sum[static_cast<size_t>(j) + static_cast<size_t>(j) * (j + 1) / 2];
    a_std[j] = std::sqrt(std::max(ajj / w_sum - aw * aw, 0.0));
  }
  const double inv_w = 1.0 / w_sum;
  Vec a_bar(d), ab_bar(d), aa_bar(tri);
  for (int j = 0; j < d; ++j) {
    if (a_std[j] == 0.0) {
      a_bar[j] = 0.0;
      ab_bar[j] = 0.0;
    } else {
      a_bar[j] = a_sum[j] * inv_w / a_std[j];
      ab_bar[j] = ab_sum[j] * inv_w / (a_std[j] * b_std);
    }
  }
  {
    size_t q = 0;
    for (int j = 0; j < d; ++j)
      for (int i = 0; i <= j; ++i, ++q)
        aa_bar[q] = (a_std[j] == 0.0 || a_std[i] == 0.0) ? 0.0 : aa_sum[q] * inv_w / (a_std[i] * a_std[j]);
  }
  const double eff_reg = p.reg_param / b_std;
  const double eff_l1 = p.elastic_net * eff_reg;
  const double eff_l2 = (1.0 - p.elastic_net) * eff_reg;
  for (int j = 0; j < d; ++j) {  // ridge term on the diagonal
    double lam = eff_l2;
    if (!p.standardization) lam = a_std[j] != 0.0 ? lam / (a_std[j] * a_std[j]) : 0.0;
    aa_bar[static_cast<size_t>(j) + static_cast<size_t>(j) * (j + 1) / 2] += lam;
  }
  const int k = p.fit_intercept ? d + 1 : d;
  auto make_ata = [&] {
    Vec m = aa_bar;
    if (p.fit_intercept) {
      m.insert(m.end(), a_bar.begin(), a_bar.end());
      m.push_back(1.0);
    }
    return m;
  };
  auto make_atb = [&] {
    Vec v = ab_bar;
    if (p.fit_intercept) v.push_back(b_bar);
    return v;
  };

  auto run_qn = [&](bool with_l1) {
    const Vec ata = make_ata(), atb = make_atb();
    CostFn cost = [&](Vec& x, Vec& grad) {
      if (p.fit_intercept) {  // the intercept slot is a function of the other coordinates; written back into x
        double s = 0.0;
        for (int j = 0; j < d; ++j) s += x[j] * a_bar[j];
        x[d] = b_bar - s;
      }
      Vec ax;
      packed_symv(k, ata, x, ax);
      const double loss = 0.5 * bb_bar - dot(x, atb) + 0.5 * dot(x, ax);
      grad.resize(k);
      for (int i = 0; i < k; ++i) grad[i] = ax[i] - atb[i];
      return loss;
    };
    L1Fn l1;
    if (with_l1) {
      l1 = [&, d](int idx) {
        if (p.fit_intercept && idx == d) return 0.0;
        if (p.standardization) return eff_l1;
        return a_std[idx] != 0.0 ? eff_l1 / a_std[idx] : 0.0;
      };
    }
    Vec x0(k, 0.0);
    if (p.fit_intercept) x0[k - 1] = b_bar;
    return quasi_newton_minimize(cost, x0, p.max_iter, p.tol, l1);
  };

  Vec sol;
  const bool use_qn = p.elastic_net != 0.0 && p.reg_param != 0.0;  // solverType Auto inside WLS
  if (use_qn) {
    QnResult r = run_qn(eff_l1 != 0.0);
    sol = r.x;
    model.objective_history = r.adjusted_values;
    model.solver_used = eff_l1 != 0.0 ? "wls-owlqn" : "wls-lbfgs";
  } else {
    Vec ata = make_ata();
    sol = make_atb();
    if (packed_cholesky_solve(ata, sol) != 0) {  // singular: Spark retries with L-BFGS on the same system
      QnResult r = run_qn(false);
      sol = r.x;
      model.objective_history = r.adjusted_values;
      model.solver_used = "wls-lbfgs-fallback";
    } else {
      model.objective_history = {0.0};
      model.solver_used = "wls-cholesky";
      // CholeskySolver also returns aaInv; WeightedLeastSquares keeps its diagonal, rescaled to original units:
      // inv_ii / (wSum * aStd_i^2) for features, inv_kk / wSum for the intercept.  O(k^3 / 3) on the host.
      Vec inv_diag;
      if (k <= 2048 && packed_spd_inverse_diag(make_ata(), k, &inv_diag)) {
        model.diag_inv_atwa.resize(k);
        for (int i = 0; i < k; ++i) {
          const double mult = (p.fit_intercept && i == k - 1) ? 1.0 : a_std[i] * a_std[i];
          model.diag_inv_atwa[i] = inv_diag[i] / (w_sum * mult);
        }
      }
    }
  }
  model.coefficients.assign(sol.begin(), sol.begin() + d);
  model.intercept = p.fit_intercept ? sol[d] * b_std : 0.0;
  for (int j = 0; j < d; ++j) model.coefficients[j] *= a_std[j] != 0.0 ? b_std / a_std[j] : 0.0;
  return model;
}

LrModel lbfgs_path_fit(const ColumnStats& st, int d, const LrParams& p, const LsqPassFn& pass) {
  if (!(st.count > 0.0)) throw SolveError{DQ4ML_ERR_EMPTY, "Training dataset is empty."};
  LrModel model;
  model.num_rows = static_cast<int64_t>(st.count);
  model.data_passes = 1;  // the summarizer pass
  const double y_mean = st.mean[d];
  const double raw_y_std = std::sqrt(st.variance[d]);
  if (raw_y_std == 0.0) {
    if (p.fit_intercept || y_mean == 0.0) {
      model.coefficients.assign(d, 0.0);
      model.intercept = y_mean;
      model.objective_history = {0.0};
      model.solver_used = "trivial";
      return model;
    }
    if (p.reg_param != 0.0)
      throw SolveError{DQ4ML_ERR_INVALID, "The standard deviation of the label is zero. Model cannot be regularized."};
  }
  const double y_std = raw_y_std > 0.0 ? raw_y_std : std::fabs(y_mean);
  Vec f_mean(st.mean.begin(), st.mean.begin() + d), f_std(d);
  for (int j = 0; j < d; ++j) f_std[j] = std::sqrt(st.variance[j]);
  const double eff_reg = p.reg_param / y_std;
  const double eff_l1 = p.elastic_net * eff_reg;
  const double eff_l2 = (1.0 - p.elastic_net) * eff_reg;

  CostFn cost = [&](Vec& c, Vec& grad) {
    double loss_sum = 0.0, w_sum = 0.0;
    Vec gsum(d, 0.0);
    pass(c, f_std, f_mean, y_std, y_mean, loss_sum, w_sum, gsum);
    double loss = loss_sum / w_sum;
    grad.resize(d);
    for (int j = 0; j < d; ++j) grad[j] = gsum[j] / w_sum;
    if (eff_l2 != 0.0) {
      double s = 0.0;
      for (int j = 0; j < d; ++j) {
        if (p.standardization) {
          s += c[j] * c[j];
          grad[j] += eff_l2 * c[j];
        } else if (f_std[j] != 0.0) {
          const double t = c[j] / (f_std[j] * f_std[j]);
          s += c[j] * t;
          grad[j] += eff_l2 * t;
        }
      }
      loss += 0.5 * s * eff_l2;
    }
    return loss;
  };
  L1Fn l1;
  const bool plain = p.elastic_net == 0.0 || eff_reg == 0.0;
  if (!plain) {
    l1 = [&](int idx) {
      if (p.standardization) return eff_l1;
      return f_std[idx] != 0.0 ? eff_l1 / f_std[idx] : 0.0;
    };
  }
  QnResult r = quasi_newton_minimize(cost, Vec(d, 0.0), p.max_iter, p.tol, l1);
  model.objective_history = r.adjusted_values;
  model.solver_used = plain ? "lbfgs" : "owlqn";
  model.data_passes += r.evaluations;
  model.coefficients = r.x;
  double s = 0.0;
  for (int j = 0; j < d; ++j) {
    model.coefficients[j] *= f_std[j] != 0.0 ? y_std / f_std[j] : 0.0;
    s += model.coefficients[j] * f_mean[j];
  }
  model.intercept = p.fit_intercept ? y_mean - s : 0.0;
  return model;
}

// LinearRegression.train with loss = huber: HuberAggregator per evaluation (label NOT standardised, features divided by
// their standard deviation), L2 penalty with regParam as given, parameters [coefficients, intercept?, sigma] started at
// all ones, LBFGSB with sigma >= Double.MinPositiveValue.
LrModel huber_path_fit(const ColumnStats& st, int d, const LrParams& p, const HuberPassFn& pass) {
  if (!(st.count > 0.0)) throw SolveError{DQ4ML_ERR_EMPTY, "Training dataset is empty."};
  LrModel model;
  model.num_rows = static_cast<int64_t>(st.count);
  model.data_passes = 1;  // the summarizer pass
  const double y_mean = st.mean[d];
  const double raw_y_std = std::sqrt(st.variance[d]);
  if (raw_y_std == 0.0) {
    if (p.fit_intercept || y_mean == 0.0) {
      model.coefficients.assign(d, 0.0);
      model.intercept = y_mean;
      model.objective_history = {0.0};
      model.solver_used = "trivial";
      return model;
    }
    if (p.reg_param != 0.0)
      throw SolveError{DQ4ML_ERR_INVALID, "The standard deviation of the label is zero. Model cannot be regularized."};
  }
  Vec f_std(d);
  for (int j = 0; j < d; ++j) f_std[j] = std::sqrt(st.variance[j]);
  const double l2 = p.reg_param;  // (1 - elasticNetParam) regParam with elasticNetParam = 0
  const int dim = d + (p.fit_intercept ? 2 : 1);

  CostFn cost = [&](Vec& c, Vec& grad) {
    double loss_sum = 0.0, w_sum = 0.0;
    Vec gsum(dim, 0.0);
    pass(c, f_std, loss_sum, w_sum, gsum);
    double loss = loss_sum / w_sum;
    grad.resize(dim);
    for (int j = 0; j < dim; ++j) grad[j] = gsum[j] / w_sum;
    if (l2 != 0.0) {
      double s = 0.0;
      for (int j = 0; j < d; ++j) {
        if (p.standardization) {
          s += c[j] * c[j];
          grad[j] += l2 * c[j];
        } else if (f_std[j] != 0.0) {
          const double t = c[j] / (f_std[j] * f_std[j]);
          s += c[j] * t;
          grad[j] += l2 * t;
        }
      }
      loss += 0.5 * s * l2;
    }
    return loss;
  };
  Vec lower(dim, -std::numeric_limits<double>::max()), upper(dim, std::numeric_limits<double>::max());
  lower[dim - 1] = std::numeric_limits<double>::denorm_min();  // Double.MinPositiveValue
  QnResult r = lbfgsb_minimize(cost, Vec(dim, 1.0), lower, upper, p.max_iter, p.tol, 10);
  model.objective_history = r.adjusted_values;
  model.solver_used = "lbfgsb";
  model.data_passes += r.evaluations;
  model.coefficients.assign(r.x.begin(), r.x.begin() + d);
  for (int j = 0; j < d; ++j) model.coefficients[j] *= f_std[j] != 0.0 ? 1.0 / f_std[j] : 0.0;
  model.intercept = p.fit_intercept ? r.x[d] : 0.0;
  model.scale = r.x[dim - 1];
  return model;
}

}  // namespace dq4ml
