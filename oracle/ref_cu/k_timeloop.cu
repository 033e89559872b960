/* TEST / MEASUREMENT INFRASTRUCTURE ONLY: the reference's rk3_g (src/timeloop.cu:34) compiled in place for sm_100a, one launch
 * per field as Timeloop::exec does (:140-180; note the argument order of the reference's call: (tendency, field)). */
#define USECUDA
#include "timeloop.cu"
#include "refcu.h"

API void refcu_rk3(const RefCuGrid* g, int substep, double* a, double* at, double dt, cudaStream_t s)
{
    dim3 gr, bl; refcu_shape(*g, g->ktot, gr, bl);
    #define RK3(S) rk3_g<double, S><<<gr, bl, 0, s>>>(a, at, dt, g->icells, g->ijcells, g->istart, g->jstart, g->kstart, g->iend, g->jend, g->kend)
    if (substep == 0) RK3(0); else if (substep == 1) RK3(1); else RK3(2);
    #undef RK3
}
