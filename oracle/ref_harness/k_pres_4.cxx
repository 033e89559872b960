/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): the reference's Pres_4 compiled in place.
 * Calls Pres_4<TF>::input/solve/output/calc_divergence (src/pres_4.cxx:256,320,533,733) the way
 * Pres_4<TF>::exec does (src/pres_4.cxx:77-141). */
#include "pres_4.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_pres4_input(void* h, double* p, const double* u, const double* v, const double* w,
                         double* ut, double* vt, double* wt, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
        W->pres4->input<false>(p, u, v, w, ut, vt, wt, gd.dzi4.data(), dt);
    else
        W->pres4->input<true>(p, u, v, w, ut, vt, wt, gd.dzi4.data(), dt);
}

API void ref_pres4_solve(void* h, double* p)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    Pres_4<double>& P = *W->pres4;
    const int jslice = 1;
    const int ns = gd.iblock*jslice*(gd.kmax+4);
    std::vector<double> tmp1(gd.ncells), tmp2(std::max(gd.ncells, 4*ns)), tmp3(std::max(gd.ncells, 4*ns));
    P.solve(p, tmp1.data(), gd.dz.data(),
            P.m1.data(), P.m2.data(), P.m3.data(), P.m4.data(), P.m5.data(), P.m6.data(), P.m7.data(),
            &tmp2[0*ns], &tmp2[1*ns], &tmp2[2*ns], &tmp2[3*ns],
            &tmp3[0*ns], &tmp3[1*ns], &tmp3[2*ns], &tmp3[3*ns],
            P.bmati.data(), P.bmatj.data(), jslice);
}

API void ref_pres4_output(void* h, double* ut, double* vt, double* wt, const double* p)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (gd.jtot == 1)
        W->pres4->output<false>(ut, vt, wt, p, gd.dzhi4.data());
    else
        W->pres4->output<true>(ut, vt, wt, p, gd.dzhi4.data());
}

API void ref_pres4_exec(void* h, double* p, const double* u, const double* v, const double* w,
                        double* ut, double* vt, double* wt, double dt)
{
    ref_pres4_input(h, p, u, v, w, ut, vt, wt, dt);
    ref_pres4_solve(h, p);
    ref_pres4_output(h, ut, vt, wt, p);
}

API double ref_pres4_divergence(void* h, const double* u, const double* v, const double* w)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    return W->pres4->calc_divergence(u, v, w, gd.dzi4.data());
}

// bands: 7*kmax doubles (m1..m7)
API void ref_pres4_coefs(void* h, double* bmati, double* bmatj, double* bands)
{
    World* W = static_cast<World*>(h);
    Pres_4<double>& P = *W->pres4;
    const int kmax = W->grid->get_grid_data().kmax;
    std::copy(P.bmati.begin(), P.bmati.end(), bmati);
    std::copy(P.bmatj.begin(), P.bmatj.end(), bmatj);
    const std::vector<double>* m[7] = {&P.m1, &P.m2, &P.m3, &P.m4, &P.m5, &P.m6, &P.m7};
    for (int n = 0; n < 7; ++n)
        std::copy(m[n]->begin(), m[n]->end(), bands + n*kmax);
}
