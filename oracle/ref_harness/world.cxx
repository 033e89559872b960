/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness). See world.h.
 * C entry points that build the reference's Grid / FFT / Pres objects and expose
 * the grid metrics computed by the reference itself (src/grid.cxx:106-189,237-368).
 */
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <stdexcept>
#include <string>

#include "world.h"

#define API extern "C" __attribute__((visibility("default")))

namespace
{
    void set_item(const std::string& k, double v)
    {
        char buf[64];
        std::snprintf(buf, sizeof(buf), "%.17g", v);
        mhh_oracle_items()[k] = buf;
    }
    void set_item(const std::string& k, int v) { mhh_oracle_items()[k] = std::to_string(v); }
}

API void* ref_world_create(int itot, int jtot, int ktot, double xsize, double ysize, double zsize,
                           int order, int igc, int jgc, int kgc, const double* z)
{
    try
    {
        auto& items = mhh_oracle_items();
        items.clear();
        set_item("grid.itot", itot); set_item("grid.jtot", jtot); set_item("grid.ktot", ktot);
        set_item("grid.xsize", xsize); set_item("grid.ysize", ysize); set_item("grid.zsize", zsize);
        items["grid.swspatialorder"] = (order == 4) ? "4" : "2";
        items["master.npx"] = "1";
        items["master.npy"] = "1";

        World* w = new World;
        w->order = order;
        w->master.start();
        w->input = new Input(w->master, "oracle");
        w->master.init(*w->input);

        w->grid = std::make_unique<Grid<double>>(w->master, *w->input);
        w->grid->set_minimum_ghost_cells(igc, jgc, kgc);   // what Advec_2i5's ctor does, src/advec_2i5.cxx:42-45
        w->grid->init();

        // Grid::create (src/grid.cxx:197-210) without the NetCDF read: z goes to [kstart, kend).
        Grid_data<double>& gd = w->grid->gd;
        for (int k = 0; k < ktot; ++k)
            gd.z[k + gd.kstart] = z[k];
        w->grid->calculate();

        // Fields: storage only. No constructor is run (it needs Boost/NetCDF-backed classes);
        // only the plain std::vector members rhoref/rhorefh and the scalar visc are used.
        void* raw = std::calloc(1, sizeof(Fields<double>));
        w->fields = reinterpret_cast<Fields<double>*>(raw);
        new (&w->fields->rhoref) std::vector<double>(gd.kcells, 1.);
        new (&w->fields->rhorefh) std::vector<double>(gd.kcells, 1.);

        w->fft = std::make_unique<FFT<double>>(w->master, *w->grid);
        w->cyclic = std::make_unique<Boundary_cyclic<double>>(w->master, *w->grid);
        w->cyclic->init();
        w->ops = std::make_unique<Field3d_operators<double>>(w->master, *w->grid, *w->fields);

        if (order == 4)
        {
            w->pres4 = std::make_unique<Pres_4<double>>(w->master, *w->grid, *w->fields, *w->fft, *w->input);
            w->pres4->init();
        }
        else
        {
            w->pres2 = std::make_unique<Pres_2<double>>(w->master, *w->grid, *w->fields, *w->fft, *w->input);
            w->pres2->init();
        }
        w->fft->load();
        w->tmp1.assign(gd.ncells, 0.);
        w->tmp2.assign(gd.ncells, 0.);
        return w;
    }
    catch (std::exception& e)
    {
        std::fprintf(stderr, "ref_world_create: %s\n", e.what());
        return nullptr;
    }
}

API void ref_world_set_rhoref(void* h, const double* rhoref, const double* rhorefh)
{
    World* w = static_cast<World*>(h);
    const int kcells = w->grid->gd.kcells;
    std::copy(rhoref, rhoref + kcells, w->fields->rhoref.begin());
    std::copy(rhorefh, rhorefh + kcells, w->fields->rhorefh.begin());
}

// Pres::set_values (src/pres_2.cxx:125, src/pres_4.cxx:179); call after ref_world_set_rhoref.
API void ref_pres_set_values(void* h)
{
    World* w = static_cast<World*>(h);
    if (w->pres2) w->pres2->set_values();
    if (w->pres4) w->pres4->set_values();
}

API void ref_world_destroy(void* h)
{
    World* w = static_cast<World*>(h);
    if (!w) return;
    w->pres2.reset();
    w->pres4.reset();
    w->ops.reset();
    w->cyclic.reset();
    w->fft.reset();
    w->fields->rhoref.~vector();
    w->fields->rhorefh.~vector();
    std::free(w->fields);
    w->grid.reset();
    delete w->input;
    delete w;
}

// out[0..15] = itot jtot ktot imax jmax kmax igc jgc kgc icells jcells kcells ijcells ncells istart/jstart/kstart...
API void ref_grid_dims(void* h, int* out)
{
    const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data();
    const int v[] = {gd.itot, gd.jtot, gd.ktot, gd.imax, gd.jmax, gd.kmax, gd.igc, gd.jgc, gd.kgc,
                     gd.icells, gd.jcells, gd.kcells, gd.ijcells, gd.ncells,
                     gd.istart, gd.iend, gd.jstart, gd.jend, gd.kstart, gd.kend};
    std::memcpy(out, v, sizeof(v));
}

API int ref_grid_array(void* h, const char* name, double* out)
{
    const Grid_data<double>& gd = static_cast<World*>(h)->grid->get_grid_data();
    const std::string n(name);
    const std::vector<double>* v = nullptr;
    if (n == "z") v = &gd.z; else if (n == "zh") v = &gd.zh;
    else if (n == "dz") v = &gd.dz; else if (n == "dzh") v = &gd.dzh;
    else if (n == "dzi") v = &gd.dzi; else if (n == "dzhi") v = &gd.dzhi;
    else if (n == "dzi4") v = &gd.dzi4; else if (n == "dzhi4") v = &gd.dzhi4;
    else if (n == "x") v = &gd.x; else if (n == "xh") v = &gd.xh;
    else if (n == "y") v = &gd.y; else if (n == "yh") v = &gd.yh;
    else if (n == "scalars")
    {
        out[0] = gd.dx; out[1] = gd.dy; out[2] = gd.dxi; out[3] = gd.dyi;
        out[4] = gd.dzhi4bot; out[5] = gd.dzhi4top;
        return 6;
    }
    if (!v) return -1;
    std::copy(v->begin(), v->end(), out);
    return (int)v->size();
}

// Boundary_cyclic::exec / exec_2d (src/boundary_cyclic.cxx:369-443,445-)
API void ref_cyclic(void* h, double* data, int edge)
{
    World* w = static_cast<World*>(h);
    const Edge e = (edge == 0) ? Edge::East_west_edge : (edge == 1) ? Edge::North_south_edge : Edge::Both_edges;
    w->cyclic->exec(data, e);
}
API void ref_cyclic_2d(void* h, double* data) { static_cast<World*>(h)->cyclic->exec_2d(data); }

// Field3d_operators (src/field3d_operators.cxx:46,69,135)
API void ref_mean_profile(void* h, double* prof, const double* fld) { static_cast<World*>(h)->ops->calc_mean_profile(prof, fld); }
API double ref_mean_2d(void* h, const double* fld) { return static_cast<World*>(h)->ops->calc_mean_2d(fld); }
API double ref_mean(void* h, const double* fld) { return static_cast<World*>(h)->ops->calc_mean(fld); }

// FFT::exec_forward / exec_backward on a ghost-free array (src/fft.cxx:588-599); backward result lands in out.
API void ref_fft_forward(void* h, double* data)
{
    World* w = static_cast<World*>(h);
    w->fft->exec_forward(data, w->tmp1.data());
}
API void ref_fft_backward(void* h, double* data, double* out)
{
    World* w = static_cast<World*>(h);
    w->fft->exec_backward(data, out);
}
