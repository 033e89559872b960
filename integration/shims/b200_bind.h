// What a maintainer adds to the reference tree (build option -DUSEB200, implies -DUSECUDA) to bind libmhh_b200.so: one
// context per run and a per-precision table of the C-ABI entry points, so that the templated reference classes
// (Advec_2<TF>, Diff_smag2<TF>, Pres_2<TF>, Timeloop<TF>; TF = double, or float with -DUSESP) call the matching symbol set
// of include/mhh_b200.h / include/mhh_b200_f32.h.  tests/test_integration_shims.py compiles these files with
// `g++ -fsyntax-only` against /root/reference/include, instantiating every shim for both precisions.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "mhh_b200.h"
#include "mhh_b200_f32.h"

template<typename TF> struct B200;

#define B200_FN(name) static constexpr auto name = mhh_##name
template<> struct B200<double>
{
    using ctx_t = mhh_ctx; using grid_desc = mhh_grid_desc;
    B200_FN(create); B200_FN(last_error); B200_FN(set_rhoref); B200_FN(pres_set_values);
    B200_FN(advec_exec); B200_FN(advec_cfl); B200_FN(advec_time_limit);
    B200_FN(smag2_exec_viscosity); B200_FN(smag2_exec); B200_FN(smag2_dnmul);
    B200_FN(pres_exec); B200_FN(pres_check_divergence); B200_FN(rk_substep); B200_FN(cyclic); B200_FN(ghost_cells_w);
    static ctx_t*& ctx() { static ctx_t* c = nullptr; return c; }
};
#undef B200_FN
#define B200_FN(name) static constexpr auto name = mhh_##name##_f32
template<> struct B200<float>
{
    using ctx_t = mhh_ctx_f32; using grid_desc = mhh_grid_desc_f32;
    B200_FN(create); B200_FN(last_error); B200_FN(set_rhoref); B200_FN(pres_set_values);
    B200_FN(advec_exec); B200_FN(advec_cfl); B200_FN(advec_time_limit);
    B200_FN(smag2_exec_viscosity); B200_FN(smag2_exec); B200_FN(smag2_dnmul);
    B200_FN(pres_exec); B200_FN(pres_check_divergence); B200_FN(rk_substep); B200_FN(cyclic); B200_FN(ghost_cells_w);
    static ctx_t*& ctx() { static ctx_t* c = nullptr; return c; }
};
#undef B200_FN

// status code -> exception, as cuda_safe_call does (include/tools.h:78-85)
template<typename TF> inline void b200_check(int err, const char* what)
{
    if (err) throw std::runtime_error(std::string(what) + ": " + B200<TF>::last_error(B200<TF>::ctx()));
}
#define B200_CALL(TF, fn, ...) b200_check<TF>(B200<TF>::fn(B200<TF>::ctx(), __VA_ARGS__), #fn)
