/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): rk3/rk4 of src/timeloop.cxx:250,288 compiled in place. */
#include "timeloop.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_rk(void* h, int rkorder, double* a, double* at, int substep, double dt)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (rkorder == 3)
        rk3<double>(a, at, substep, dt, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.ncells);
    else
        rk4<double>(a, at, substep, dt, gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend,
                gd.icells, gd.ijcells, gd.ncells);
}

API double ref_rk_subdt(int rkorder, double dt, int substep)
{
    return (rkorder == 3) ? rk3subdt<double>(dt, substep) : rk4subdt<double>(dt, substep);
}

// Timeloop::save / load (src/timeloop.cxx:445-525) on a Timeloop built by the reference's own constructor from the in-memory
// [time] block (starttime 0, endtime, savetime, dt, rkorder); files "time.%07d" in the current working directory.
namespace
{
    Timeloop<double>* make_timeloop(World* W)
    {
        auto& it = mhh_oracle_items();
        it["time.starttime"] = "0"; it["time.endtime"] = "3600"; it["time.savetime"] = "3600";
        it["time.dt"] = "1"; it["time.rkorder"] = "3";
        alignas(64) static char soil_dummy[4096];       // only the reference to it is stored
        return new Timeloop<double>(W->master, *W->grid, *reinterpret_cast<Soil_grid<double>*>(soil_dummy), *W->fields, *W->input,
                                    Sim_mode::Run);
    }
}

API void ref_time_save(void* h, int iotime, unsigned long itime, unsigned long idt, int iteration)
{
    World* W = static_cast<World*>(h);
    Timeloop<double>* tl = make_timeloop(W);
    tl->save(iotime, itime, idt, iteration);
    delete tl;
}

// out[0..2] = itime, idt, iteration read back by the reference's own loader
API void ref_time_load(void* h, int iotime, unsigned long* out)
{
    World* W = static_cast<World*>(h);
    Timeloop<double>* tl = make_timeloop(W);
    tl->load(iotime);
    out[0] = tl->itime; out[1] = tl->idt; out[2] = static_cast<unsigned long>(tl->iteration);
    delete tl;
}
