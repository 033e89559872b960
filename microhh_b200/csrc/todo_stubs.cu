// Temporary: schemes not built yet return MHH_ERR_UNSUPPORTED (loud, never a fallback).
#include "ctx.h"
namespace mhh
{
int advec2i5_exec(Ctx* c, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*, const int*) { c->last_error = "advec_2i5 not built"; return MHH_ERR_UNSUPPORTED; }
int advec2i5_cfl_bits(Ctx* c, const double*, const double*, const double*, unsigned long long*) { return MHH_ERR_UNSUPPORTED; }
int advec4_exec(Ctx* c, const double*, const double*, const double*, double*, double*, double*, int, const double* const*, double* const*) { c->last_error = "advec_4 not built"; return MHH_ERR_UNSUPPORTED; }
int advec4_cfl_bits(Ctx* c, const double*, const double*, const double*, unsigned long long*) { return MHH_ERR_UNSUPPORTED; }
int diff4_exec(Ctx* c, const double*, const double*, const double*, double*, double*, double*, double, int, const double* const*, double* const*, const double*) { c->last_error = "diff_4 not built"; return MHH_ERR_UNSUPPORTED; }
int pres4_set_values(Ctx* c) { c->last_error = "pres_4 not built"; return MHH_ERR_UNSUPPORTED; }
int pres4_input(Ctx*, const double*, const double*, const double*, const double*, const double*, double*, double, double*) { return MHH_ERR_UNSUPPORTED; }
int pres4_solve_spectral(Ctx*, double*) { return MHH_ERR_UNSUPPORTED; }
int pres4_output(Ctx*, double*, double*, double*, double*) { return MHH_ERR_UNSUPPORTED; }
int pres4_divergence_bits(Ctx*, const double*, const double*, const double*, unsigned long long*) { return MHH_ERR_UNSUPPORTED; }
}
