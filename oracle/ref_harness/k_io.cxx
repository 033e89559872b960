/* TEST INFRASTRUCTURE ONLY.  The reference's restart-file writers / readers compiled in place and driven on raw arrays:
 * Field3d_io::save_field3d / load_field3d (src/field3d_io.cxx:671-751, serial build), Grid::save_grid
 * (src/grid_serial.cxx:46-111) and — in k_timeloop.cxx — Timeloop::save / load (src/timeloop.cxx:445-525).
 * Files land in the current working directory, as with `microhh init`. */
#include "field3d_io.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API int ref_save_field3d(void* h, double* data, const char* filename, double offset)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    Field3d_io<double> io(W->master, *W->grid);
    std::vector<double> t1(gd.ncells), t2(gd.ncells);
    return io.save_field3d(data, t1.data(), t2.data(), filename, offset, gd.kstart, gd.kend);
}

API int ref_load_field3d(void* h, double* data, const char* filename, double offset)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    Field3d_io<double> io(W->master, *W->grid);
    std::vector<double> t1(gd.ncells), t2(gd.ncells);
    return io.load_field3d(data, t1.data(), t2.data(), filename, offset, gd.kstart, gd.kend);
}

API void ref_save_grid(void* h)
{
    World* W = static_cast<World*>(h);
    W->grid->save_grid();
}
