/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness): vertical ghost-cell kernels of src/boundary.cxx:687-907
 * compiled in place. bc: 0 Dirichlet, 1 Neumann, 2 Flux, 3 Ustar, 4 Off (enum Boundary_type, include/boundary.h:51). */
#include "boundary.cxx"
#include "world.h"
#define API extern "C" __attribute__((visibility("default")))

API void ref_ghost_cells(void* h, double* a, int bcbot, int bctop, double* abot, double* atop, double* agradbot, double* agradtop)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    const Boundary_type bb = static_cast<Boundary_type>(bcbot);
    const Boundary_type bt = static_cast<Boundary_type>(bctop);
    if (W->order == 2)
    {
        calc_ghost_cells_bot_2nd<double>(a, gd.dzh.data(), bb, abot, agradbot, gd.kstart, gd.icells, gd.jcells, gd.ijcells);
        calc_ghost_cells_top_2nd<double>(a, gd.dzh.data(), bt, atop, agradtop, gd.kend, gd.icells, gd.jcells, gd.ijcells);
    }
    else
    {
        calc_ghost_cells_bot_4th<double>(a, gd.z.data(), bb, abot, agradbot, gd.kstart, gd.icells, gd.jcells, gd.ijcells);
        calc_ghost_cells_top_4th<double>(a, gd.z.data(), bt, atop, agradtop, gd.kend, gd.icells, gd.jcells, gd.ijcells);
    }
}

// type: 0 Normal, 1 Conservation (enum Boundary_w_type). 4th order only, as in Boundary::set_ghost_cells_w :986.
API void ref_ghost_cells_w(void* h, double* w, int type)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    if (W->order != 4) return;
    if (type == 0)
    {
        calc_ghost_cells_botw_4th<double>(w, gd.kstart, gd.icells, gd.jcells, gd.ijcells);
        calc_ghost_cells_topw_4th<double>(w, gd.kend, gd.icells, gd.jcells, gd.ijcells);
    }
    else
    {
        calc_ghost_cells_botw_cons_4th<double>(w, gd.kstart, gd.icells, gd.jcells, gd.ijcells);
        calc_ghost_cells_topw_cons_4th<double>(w, gd.kend, gd.icells, gd.jcells, gd.ijcells);
    }
}

// set_bc<TF,true> (src/boundary.cxx:50-91) as used by Boundary::set_values (:598-682).
API void ref_set_bc(void* h, double* a, double* agrad, double* aflux, int sw, double aval, double visc, double offset)
{
    World* W = static_cast<World*>(h);
    const Grid_data<double>& gd = W->grid->get_grid_data();
    set_bc<double, true>(a, agrad, aflux, static_cast<Boundary_type>(sw), aval, visc, offset, gd.icells, gd.jcells);
}
