/* TEST INFRASTRUCTURE ONLY.  The reference's profile statistics compiled in place from src/stats.cxx (NetCDF replaced by
 * oracle/shim/netcdf.h): calc_mean :264, calc_moment :342, calc_cov :368 over the default mask (every point, nmask = itot*jtot). */
#include "stats.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

namespace
{
    struct AllPoints
    {
        std::vector<unsigned int> mask;
        std::vector<int> nmask;
        explicit AllPoints(const Grid_data<double>& gd) : mask(gd.ncells, 1u), nmask(gd.kcells, gd.itot * gd.jtot) {}
    };
}

API void ref_stats_mean(void* h, double* prof, const double* fld)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    AllPoints m(gd);
    calc_mean<double>(prof, fld, m.mask.data(), 1u, m.nmask.data(), gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                      gd.icells, gd.ijcells);
}

API void ref_stats_moment(void* h, double* prof, const double* fld, const double* fld_mean, double offset, int power)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    AllPoints m(gd);
    calc_moment<double>(prof, fld, fld_mean, offset, m.mask.data(), 1u, m.nmask.data(), power,
                        gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}

API void ref_stats_cov(void* h, double* prof, const double* fld1, const double* mean1, double off1, int pow1,
                       const double* fld2, const double* mean2, double off2, int pow2)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    AllPoints m(gd);
    calc_cov<double>(prof, fld1, mean1, off1, pow1, fld2, mean2, off2, pow2, m.mask.data(), 1u, m.nmask.data(),
                     gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend, gd.icells, gd.ijcells);
}
