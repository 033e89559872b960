/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness).
 *
 * "World" = the handful of reference objects the hot-path arithmetic needs
 * (Master, Grid, FFT, Boundary_cyclic, Field3d_operators, Pres_2, Pres_4),
 * constructed through the reference's OWN constructors from the reference's own
 * sources compiled in place (never copied).  Built with -fno-access-control so
 * the harness may call private members such as Pres_2<TF>::input/solve/output
 * (src/pres_2.cxx:156,267,365) and Grid<TF>::calculate (src/grid.cxx:237).
 */
#ifndef MHH_ORACLE_WORLD_H
#define MHH_ORACLE_WORLD_H

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "master.h"
#include "input.h"
#include "grid.h"
#include "fields.h"
#include "fft.h"
#include "boundary_cyclic.h"
#include "field3d_operators.h"
#include "pres_2.h"
#include "pres_4.h"

// key = "block.item", value = textual value; consumed by input_stub.cxx
std::map<std::string, std::string>& mhh_oracle_items();

struct World
{
    Master master;
    Input* input;
    std::unique_ptr<Grid<double>> grid;
    Fields<double>* fields;  // zero-filled storage; only rhoref/rhorefh/visc are ever touched
    std::unique_ptr<FFT<double>> fft;
    std::unique_ptr<Boundary_cyclic<double>> cyclic;
    std::unique_ptr<Field3d_operators<double>> ops;
    std::unique_ptr<Pres_2<double>> pres2;
    std::unique_ptr<Pres_4<double>> pres4;
    std::vector<double> tmp1, tmp2;
    int order;
};

#endif
