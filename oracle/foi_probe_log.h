/* oracle/foi_probe_log.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * The record an instrumented build of the reference's Community.cpp leaves for every force-of-infection it computes
 * (see foi_probe.h). Shared by the instrumented object and by ref_driver.cpp (mode `foi`). */
#ifndef CABM_FOI_PROBE_LOG_H
#define CABM_FOI_PROBE_LOG_H
#include <cmath>
#include <vector>
struct cabm_probe_rec { int line; int loc; double x; double weight; };       /* x = the argument of exp(), i.e. -hazard */
inline std::vector<cabm_probe_rec> cabm_probe_log;
static inline double cabm_probe_exp(double x, int line, int loc, double weight) {
    cabm_probe_log.push_back(cabm_probe_rec{ line, loc, x, weight });
    return std::exp(x);
}
#endif
