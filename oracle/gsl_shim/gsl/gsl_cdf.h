/* oracle/gsl_shim/gsl/gsl_cdf.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 * Only gsl_cdf_logistic_Pinv is used by the reference (Parameters.h:631). */
#ifndef GSL_SHIM_CDF_H
#define GSL_SHIM_CDF_H
#include <cmath>
#include <limits>
static inline double gsl_cdf_logistic_Pinv(const double P, const double a) {
    if (P == 1.0) return  std::numeric_limits<double>::infinity();
    if (P == 0.0) return -std::numeric_limits<double>::infinity();
    return a * std::log(P / (1.0 - P));   /* quantile of the logistic distribution with scale a */
}
#endif
