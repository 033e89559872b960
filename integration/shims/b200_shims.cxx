// The guarded member bodies of a -DUSEB200 build (see b200_bind.h and INTEGRATION.md): the same members the reference's
// .cu files define under #ifdef USECUDA (src/advec_2.cu:140-226, src/diff_smag2.cu:553-790, src/pres_2.cu:253-360,
// src/timeloop.cu:127-183), each gathering the device pointers of Fields and calling one extern "C" entry.
#define USECUDA
#include "master.h"
#include "grid.h"
#include "fields.h"
#include "field3d.h"
#include "stats.h"
#include "thermo.h"
#include "boundary.h"
#include "advec_2.h"
#include "diff_smag2.h"
#include "pres_2.h"
#include "timeloop.h"
#include "soil_grid.h"
#include "soil_field3d.h"
#include "constants.h"

#include <cuda_runtime_api.h>

#include "b200_bind.h"

// ---- once per run: Model::prepare_gpu (src/model.cxx:561-579) ----------------------------------------------------------
template<typename TF>
void b200_create(Grid<TF>& grid, Fields<TF>& fields)
{
    const Grid_data<TF>& gd = grid.get_grid_data();
    typename B200<TF>::grid_desc d{gd.itot, gd.jtot, gd.ktot, gd.igc, gd.jgc, gd.kgc, gd.xsize, gd.ysize, gd.zsize,
                                   grid.get_spatial_order() == Grid_order::Second ? 2 : 4, /*npy*/ 1, /*rank_y*/ 0};
    if (B200<TF>::create(&d, &gd.z[gd.kstart], &B200<TF>::ctx())) throw std::runtime_error("mhh_create failed");
    B200_CALL(TF, set_rhoref, fields.rhoref.data(), fields.rhorefh.data());
    b200_check<TF>(B200<TF>::pres_set_values(B200<TF>::ctx()), "pres_set_values");
}

namespace
{
    // the scalars in std::map order, as the reference's loops over fields.st visit them
    template<typename TF> struct Scalars
    {
        std::vector<const TF*> s, fb, ft; std::vector<TF*> st; std::vector<TF> visc;
        explicit Scalars(Fields<TF>& f)
        {
            for (auto& it : f.st)
            {
                st.push_back(it.second->fld_g); s.push_back(f.sp.at(it.first)->fld_g);
                fb.push_back(f.sp.at(it.first)->flux_bot_g); ft.push_back(f.sp.at(it.first)->flux_top_g);
                visc.push_back(f.sp.at(it.first)->visc);
            }
        }
    };
}

// ---- Advec_2 (src/advec_2.cu:140-226) -------------------------------------------------------------------------------------
template<typename TF>
void Advec_2<TF>::exec(Stats<TF>& stats)
{
    Scalars<TF> sc(fields);
    B200_CALL(TF, advec_exec, MHH_ADVEC_2, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g,
              fields.mt.at("u")->fld_g, fields.mt.at("v")->fld_g, fields.mt.at("w")->fld_g,
              (int)sc.s.size(), sc.s.data(), sc.st.data(), nullptr);
    cudaDeviceSynchronize();
    stats.calc_tend(*fields.mt.at("u"), tend_name);
    stats.calc_tend(*fields.mt.at("v"), tend_name);
    stats.calc_tend(*fields.mt.at("w"), tend_name);
    for (auto it : fields.st)
        stats.calc_tend(*it.second, tend_name);
}

template<typename TF>
double Advec_2<TF>::get_cfl(const double dt)
{
    TF cfl;
    B200_CALL(TF, advec_cfl, MHH_ADVEC_2, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g, dt, &cfl);
    return static_cast<double>(cfl);
}

template<typename TF>
unsigned long Advec_2<TF>::get_time_limit(unsigned long idt, double dt)
{
    unsigned long lim;
    B200_CALL(TF, advec_time_limit, MHH_ADVEC_2, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g,
              idt, dt, static_cast<TF>(cflmax), &lim);
    return lim;
}

// ---- Diff_smag2 (src/diff_smag2.cu:553-880) -------------------------------------------------------------------------------
template<typename TF>
void Diff_smag2<TF>::exec_viscosity(Thermo<TF>& thermo)
{
    const bool sm = boundary.get_switch() != "default";
    TF* N2 = nullptr;
    auto tmp = fields.get_tmp_g();
    if (thermo.get_switch() != "0")
    {
        thermo.get_thermo_field_g(*tmp, "N2", false);
        N2 = tmp->fld_g;
    }
    B200_CALL(TF, smag2_exec_viscosity, sm ? 1 : 0, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g, N2,
              sm ? boundary.get_dudz_g() : nullptr, sm ? boundary.get_dvdz_g() : nullptr,
              (sm && N2) ? boundary.get_dbdz_g() : nullptr, sm ? boundary.get_z0m_g() : nullptr,
              static_cast<TF>(cs), static_cast<TF>(tPr), fields.visc, fields.sd.at("evisc")->fld_g);
    fields.release_tmp_g(tmp);
}

template<typename TF>
void Diff_smag2<TF>::exec(Stats<TF>& stats)
{
    Scalars<TF> sc(fields);
    B200_CALL(TF, smag2_exec, boundary.get_switch() != "default" ? 1 : 0,
              fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g, fields.sd.at("evisc")->fld_g,
              fields.mt.at("u")->fld_g, fields.mt.at("v")->fld_g, fields.mt.at("w")->fld_g,
              fields.mp.at("u")->flux_bot_g, fields.mp.at("u")->flux_top_g, fields.mp.at("v")->flux_bot_g, fields.mp.at("v")->flux_top_g,
              fields.visc, (int)sc.s.size(), sc.s.data(), sc.st.data(), sc.fb.data(), sc.ft.data(), sc.visc.data(), static_cast<TF>(tPr));
    cudaDeviceSynchronize();
    stats.calc_tend(*fields.mt.at("u"), tend_name);
    stats.calc_tend(*fields.mt.at("v"), tend_name);
    stats.calc_tend(*fields.mt.at("w"), tend_name);
    for (auto it : fields.st)
        stats.calc_tend(*it.second, tend_name);
}

template<typename TF>
double Diff_smag2<TF>::get_dn(const double dt)
{
    TF dnmul;
    B200_CALL(TF, smag2_dnmul, fields.sd.at("evisc")->fld_g, static_cast<TF>(tPr), &dnmul);
    return static_cast<double>(dnmul) * dt;
}

template<typename TF>
unsigned long Diff_smag2<TF>::get_time_limit(const unsigned long idt, const double dt)
{
    TF dnmul;
    B200_CALL(TF, smag2_dnmul, fields.sd.at("evisc")->fld_g, static_cast<TF>(tPr), &dnmul);
    return idt * dnmax / (dt * std::max(static_cast<double>(dnmul), Constants::dsmall));
}

// ---- Pres_2 (src/pres_2.cu:253-360) ------------------------------------------------------------------------------------------
template<typename TF>
void Pres_2<TF>::exec(double dt, Stats<TF>& stats)
{
    B200_CALL(TF, pres_exec, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g,
              fields.mt.at("u")->fld_g, fields.mt.at("v")->fld_g, fields.mt.at("w")->fld_g, fields.sd.at("p")->fld_g, dt);
    cudaDeviceSynchronize();
    stats.calc_tend(*fields.mt.at("u"), tend_name);
    stats.calc_tend(*fields.mt.at("v"), tend_name);
    stats.calc_tend(*fields.mt.at("w"), tend_name);
}

template<typename TF>
TF Pres_2<TF>::check_divergence()
{
    TF div;
    B200_CALL(TF, pres_check_divergence, fields.mp.at("u")->fld_g, fields.mp.at("v")->fld_g, fields.mp.at("w")->fld_g, &div);
    return div;
}

// ---- Timeloop (src/timeloop.cu:127-183) --------------------------------------------------------------------------------------
template<typename TF>
void Timeloop<TF>::exec()
{
    std::vector<TF*> a, at;
    for (auto& f : fields.at)
    {
        at.push_back(f.second->fld_g);
        a.push_back(fields.ap.at(f.first)->fld_g);
    }
    B200_CALL(TF, rk_substep, rkorder, substep, dt, (int)a.size(), a.data(), at.data());
    substep = (substep + 1) % (rkorder == 3 ? 3 : 5);
}

template void b200_create<double>(Grid<double>&, Fields<double>&);
template void b200_create<float>(Grid<float>&, Fields<float>&);
template class Advec_2<double>;
template class Advec_2<float>;
template class Diff_smag2<double>;
template class Diff_smag2<float>;
template class Pres_2<double>;
template class Pres_2<float>;
template class Timeloop<double>;
template class Timeloop<float>;
