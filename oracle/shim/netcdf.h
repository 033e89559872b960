/* TEST INFRASTRUCTURE ONLY (oracle). Empty stand-in for <netcdf.h>: the reference
 * headers (include/netcdf_interface.h:28) include it, but no NetCDF function is
 * reachable from the hot-path arithmetic the oracle calls. */
#ifndef MHH_ORACLE_NETCDF_SHIM_H
#define MHH_ORACLE_NETCDF_SHIM_H
#include <stddef.h>
typedef int nc_type;
#define NC_UNLIMITED 0L
#define NC_NOERR 0
#endif
