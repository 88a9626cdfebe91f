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

}  // namespace

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

}  // namespace dq4ml
