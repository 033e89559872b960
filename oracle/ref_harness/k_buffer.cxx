/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): calc_buffer of src/buffer.cxx:37-58 compiled in place, with the
 * bufferkstart/bufferkstarth search restated from Buffer::create (src/buffer.cxx:107-125; that member itself reads
 * NetCDF input and cannot run here). */
#include "buffer.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API int ref_buffer_kstart(void* h, double zstart, int half)
{
    const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data();
    int ks = gd.kstart;
    for (int k = gd.kstart; k < gd.kend; ++k)
        if ((half ? gd.zh[k] : gd.z[k]) < zstart) ++ks;
    return ks;
}

API void ref_buffer(void* h, double* at, const double* a, const double* abuf, double zstart, double sigma, double beta, int is_w)
{
    const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data();
    calc_buffer<double>(at, a, abuf, is_w ? gd.zh.data() : gd.z.data(), zstart, gd.zsize, beta, sigma,
            gd.istart, gd.iend, gd.icells, gd.jstart, gd.jend, gd.ijcells, ref_buffer_kstart(h, zstart, is_w), gd.kend);
}
