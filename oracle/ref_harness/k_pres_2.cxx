/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the reference's Pres_2 compiled in place.
 * Calls Pres_2<TF>::input/solve/output/calc_divergence (src/pres_2.cxx:156,267,365,391). */
#include "pres_2.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_pres2_input(void* h, double* p, const double* u, const double* v, const double* w,
                         double* ut, double* vt, double* wt, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    W->pres2->input(p, u, v, w, ut, vt, wt, gd.dzi.data(), W->fields->rhoref.data(), W->fields->rhorefh.data(), dt);
}

API void ref_pres2_solve(void* h, double* p)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    W->pres2->solve(p, W->tmp1.data(), W->tmp2.data(), gd.dz.data(), W->fields->rhoref.data());
}

API void ref_pres2_output(void* h, double* ut, double* vt, double* wt, const double* p)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    W->pres2->output(ut, vt, wt, p, gd.dzhi.data());
}

// Pres_2<TF>::exec (src/pres_2.cxx:66-96) minus the stats calls.
API void ref_pres2_exec(void* h, double* p, const double* u, const double* v, const double* w,
                        double* ut, double* vt, double* wt, double dt)
{
    ref_pres2_input(h, p, u, v, w, ut, vt, wt, dt);
    ref_pres2_solve(h, p);
    ref_pres2_output(h, ut, vt, wt, p);
}

API double ref_pres2_divergence(void* h, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return W->pres2->calc_divergence(u, v, w, gd.dzi.data(), W->fields->rhoref.data(), W->fields->rhorefh.data());
}

API void ref_pres2_coefs(void* h, double* bmati, double* bmatj, double* a, double* c)
{
    World* W = static_cast<World*>(h);
    std::copy(W->pres2->bmati.begin(), W->pres2->bmati.end(), bmati);
    std::copy(W->pres2->bmatj.begin(), W->pres2->bmatj.end(), bmatj);
    std::copy(W->pres2->a.begin(), W->pres2->a.end(), a);
    std::copy(W->pres2->c.begin(), W->pres2->c.end(), c);
}

// tdma (src/pres_2.cxx:202-263) on caller-supplied arrays.
API void ref_pres2_tdma(double* a, double* b, double* c, double* p, double* work2d, double* work3d,
                        int iblock, int jblock, int kmax)
{
    tdma<double>(a, b, c, p, work2d, work3d, iblock, jblock, kmax);
}
