/* TEST INFRASTRUCTURE ONLY (oracle). Empty stand-in for <netcdf.h>: the reference
 * headers (include/netcdf_interface.h:28) include it, but no NetCDF function is
 * reachable from the hot-path arithmetic the oracle calls. */
#ifndef MHH_ORACLE_NETCDF_SHIM_H
#define MHH_ORACLE_NETCDF_SHIM_H
#include <stddef.h>
typedef int nc_type;
#define NC_UNLIMITED 0L
#define NC_NOERR 0
/* the library's default fill values (netcdf.h), used by src/stats.cxx:52-53 as "no data" markers */
#define NC_FILL_DOUBLE (9.9692099683868690e+36)
#define NC_FILL_FLOAT (9.9692099683868690e+36f)
#endif
