#include "gram.hpp"
namespace dq4ml {
int gram_wls_moments(dq4ml_ctx*, const DeviceCols& cols, const double*, const DeviceDq&, int64_t, double*) {
  set_error("WLS moments for %d > 8 feature columns: Gram kernel not built yet", cols.n);
  return DQ4ML_ERR_UNSUPPORTED;
}
}
