/* oracle/foi_probe.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Compile-time instrumentation of the UNMODIFIED reference Community.cpp ("Reference-B" of SURVEY.md 8c): the file is
 * compiled where it lies with `-include foi_probe.h`. Every per-location exposure probability of Community::tick is
 * written as `const double T = 1.0 - exp(-hazard);` inside a loop over `[loc, infectious_groups]` with `infectious_weight`
 * in scope (Community.cpp:846,868,903,923,937,965); those are the only uses of exp() in the file. The macro below routes
 * them through cabm_probe_exp, which logs (source line, location id, -hazard, infectious weight) and returns std::exp.
 * No reference source is edited or copied; the arithmetic is unchanged.
 *
 * Everything Community.cpp includes is included here first, so that the macro cannot touch declarations or inline
 * functions of the standard library, the GSL stand-in or the reference's own headers. */
#ifndef CABM_FOI_PROBE_H
#define CABM_FOI_PROBE_H
#include <cstdlib>
#include <cstring>
#include <climits>
#include <iostream>
#include <string>
#include <iterator>
#include <fstream>
#include <sstream>
#include <assert.h>
#include <math.h>
#include <cmath>
#include <complex>
#include <random>
#include <algorithm>
#include <numeric>
#include <gsl/gsl_rng.h>
#include <gsl/gsl_randist.h>
#include "Person.h"
#include "Location.h"
#include "Community.h"
#include "Parameters.h"
#include "Vac_Campaign.h"
#include "foi_probe_log.h"
#define exp(x) cabm_probe_exp((x), __LINE__, (int) loc->getID(), (double) infectious_weight)
#endif
