// Monin-Obukhov surface layer, Boundary_surface<TF>::exec (src/boundary_surface.cxx:827-981 with
// include/boundary_surface_kernels.h and include/monin_obukhov.h): total wind at the first level, Obukhov length from the
// z/L lookup table (swconstantz0=true), u*, the surface fluxes / values / gradients of u, v and the scalars, and the MO
// gradients dudz, dvdz, dbdz that Diff_smag2's first row consumes.  Everything is 2-D (ijcells threads), a few dozen
// flops and four transcendentals per cell: negligible next to the 3-D passes, so the kernels are written for clarity, one
// per reference loop nest, with the cyclic halo fills of the reference in between.  The lookup table is built on the host
// exactly as prepare_lut does (double arithmetic, float storage).
//
// Not bit-exact by construction: log/pow/exp come from the CUDA math library instead of glibc (<= 2 ulp each); parity is
// checked at 1e-12 relative in tests/test_gpu_surface.py.
#include "sys.h"
#include <cmath>
#include <vector>

#include "ctx.h"

namespace mhh
{
int cyclic_exec_2d(Ctx*, double*);

namespace
{
    constexpr double kKappa = 0.4, kGrav = 9.81, kDbig = 1e9, kDhugeS = 1e30;
    constexpr double kZLmax = 10., kZLmin = -1.e4;
    constexpr int NLUT = 10000;                      // nzL_lut, include/boundary.h:55
    enum { BC_DIRICHLET = 0, BC_NEUMANN = 1, BC_FLUX = 2, BC_USTAR = 3 };

    // ---- include/monin_obukhov.h:41-149 (host + device: the table is built on the host) ----
    __host__ __device__ inline double phim_unstable(double zeta) { return pow(1. + 3.6 * pow(fabs(zeta), 2. / 3.), -1. / 2.); }
    __host__ __device__ inline double phih_unstable(double zeta) { return pow(1. + 7.9 * pow(fabs(zeta), 2. / 3.), -1. / 2.); }
    __host__ __device__ inline double phim(double zeta) { return (zeta <= 0.) ? phim_unstable(zeta) : 1. + 5. * zeta; }
    __host__ __device__ inline double phih(double zeta)
    { if (zeta <= 0.) return phih_unstable(zeta); const double a = 1. + 4. * zeta; return a * a; }
    __host__ __device__ inline double psim_unstable(double zeta) { return 3. * log((1. + 1. / phim_unstable(zeta)) / 2.); }
    __host__ __device__ inline double psih_unstable(double zeta) { return 3. * log((1. + 1. / phih_unstable(zeta)) / 2.); }
    __host__ __device__ inline double psim_stable(double zeta)
    {
        const double a = 1., b = 2. / 3., c = 5., d = 0.35;
        return -b * (zeta - (c / d)) * exp(-d * zeta) - a * zeta - (b * c) / d;
    }
    __host__ __device__ inline double psih_stable(double zeta)
    {
        const double a = 1., b = 2. / 3., c = 5., d = 0.35;
        return -b * (zeta - (c / d)) * exp(-d * zeta) - pow(1. + b * a * zeta, 1.5) - (b * c) / d + 1.;
    }
    __host__ __device__ inline double most_fm(double zsl, double z0m, double L)
    {
        return (L <= 0.) ? kKappa / (log(zsl / z0m) - psim_unstable(zsl / L) + psim_unstable(z0m / L))
                         : kKappa / (log(zsl / z0m) - psim_stable(zsl / L) + psim_stable(z0m / L));
    }
    __host__ __device__ inline double most_fh(double zsl, double z0h, double L)
    {
        return (L <= 0.) ? kKappa / (log(zsl / z0h) - psih_unstable(zsl / L) + psih_unstable(z0h / L))
                         : kKappa / (log(zsl / z0h) - psih_stable(zsl / L) + psih_stable(z0h / L));
    }

    // ---- iterative Obukhov-length solvers for heterogeneous z0 (swconstantz0=false), include/boundary_surface_kernels.h:
    // calc_obuk_noslip_flux_iterative :287-375 and calc_obuk_noslip_dirichlet_iterative :377-470.  Newton iteration on
    // f(L) = zsl/L - Ri-like term with a +-0.1 % finite-difference slope, one restart with a safe start value, the result
    // limited to the z/L range of the lookup solver.  FLUX: the first form (bfluxbot), else the second (db, z0h).
    template<bool FLUX>
    __host__ __device__ inline double obuk_iterative(double L, const double du, double bf, const double zsl, const double z0m, const double z0h)
    {
        constexpr double kDsmallS = 1.e-9, Lmax = 1.e16;
        const double L_min_stable = zsl / kZLmax;
        if (!FLUX)
        {
            // no zero crossing for large Ri: the reference returns the stable limit
            const double Ri = kKappa * bf * zsl / (du * du);
            if (Ri > 0.13) return L_min_stable;
        }
        if (bf >= 0.) bf = fmax(kDsmallS, bf); else bf = fmin(-kDsmallS, bf);
        auto fx_of = [&](const double l) -> double
        {
            if (FLUX) { const double a = du * most_fm(zsl, z0m, l); return zsl / l + kKappa * zsl * bf / (a * a * a); }
            const double a = du * most_fm(zsl, z0m, l);
            return zsl / l - kKappa * zsl * bf * most_fh(zsl, z0h, l) / (a * a);
        };
        int m = 0, nlim = 10;
        double L0;
        while (m <= 1)
        {
            // the stability changed sign, or the last attempt did not converge: restart from a safe value
            const bool reset = FLUX ? (m == 1 || L * bf >= 0.) : (m == 1 || L * bf <= 0.);
            if (reset)
            {
                nlim = 200;
                if (FLUX) L = (bf >= 0.) ? -kDsmallS : L_min_stable;
                else      L = (bf >= 0.) ? L_min_stable : -kDsmallS;
            }
            if (FLUX) L0 = (bf >= 0.) ? -kDhugeS : kDhugeS;
            else      L0 = (bf >= 0.) ? kDhugeS : -kDhugeS;
            int n = 0;
            while (fabs((L - L0) / L0) > 0.001 && n < nlim && fabs(L) < Lmax)
            {
                L0 = L;
                const double fx = fx_of(L);
                const double Lstart = L - 0.001 * L, Lend = L + 0.001 * L;
                const double fxdif = (fx_of(Lend) - fx_of(Lstart)) / (Lend - Lstart);
                if (fabs(fxdif) < 1e-16) break;
                L = L - fx / fxdif;
                if (L >= 0. && L < L_min_stable) L = L_min_stable;
                ++n;
            }
            if (n < nlim && fabs(L) < Lmax) break;
            ++m;
            nlim = 200;
        }
        if (m > 1)
        {
            if (fabs(L) > Lmax) L = (FLUX ? -copysign(1., bf) : copysign(1., bf)) * Lmax;
            else L = L_min_stable;
        }
        return zsl / fmin(fmax(zsl / L, kZLmin), kZLmax);
    }

    // calc_z0_charnock, src/boundary_surface.cxx:342-365 (interior; the cyclic fills follow)
    __global__ void charnock_kernel(double* __restrict__ z0m, double* __restrict__ z0h, const double* __restrict__ ustar,
                                    const double alpha_m, const double alpha_ch, const double alpha_h, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int ij = i + j * gd.icells;
        const double visc = 1.5e-5, gi = 1. / kGrav, min_ustar = 1e-8;
        const double ustar_lim = fmax(ustar[ij], min_ustar);
        z0m[ij] = alpha_m * visc / ustar_lim + alpha_ch * (ustar_lim * ustar_lim) * gi;
        z0h[ij] = alpha_h * visc / ustar_lim;
    }

    // find_zL, include/boundary_surface_kernels.h:249-263 (float comparisons, the search resumes at the cell's last index)
    __device__ inline double find_zL(const float* __restrict__ zL, const float* __restrict__ f, int& n, const float Ri)
    {
        if ((f[n] - Ri) > 0.f)
            while (n > 0 && (f[n - 1] - Ri) > 0.f) --n;
        else
            while ((f[n] - Ri) < 0.f && n < (NLUT - 1)) ++n;
        return (n == 0 || n == NLUT - 1) ? (double)zL[n]
                                         : zL[n - 1] + (Ri - f[n - 1]) / (f[n] - f[n - 1]) * (zL[n] - zL[n - 1]);
    }

    // calc_dutot, include/boundary_surface_kernels.h:140-188 (interior; the cyclic fill follows)
    __global__ void dutot_kernel(double* __restrict__ dutot, const double* __restrict__ u, const double* __restrict__ v,
                                 const double* __restrict__ ubot, const double* __restrict__ vbot, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int ii = 1, ii2 = 2, jj = gd.icells, jj2 = 2 * gd.icells;
        const int ij = i + j * jj, ijk = ij + gd.kstart * gd.ijcells;
        const double u_filtered = (1. / 9) *
            ( 0.5 * u[ijk - ii - jj] + u[ijk - jj] + u[ijk + ii - jj] + 0.5 * u[ijk + ii2 - jj]
            + 0.5 * u[ijk - ii     ] + u[ijk     ] + u[ijk + ii     ] + 0.5 * u[ijk + ii2     ]
            + 0.5 * u[ijk - ii + jj] + u[ijk + jj] + u[ijk + ii + jj] + 0.5 * u[ijk + ii2 + jj] );
        const double v_filtered = (1. / 9) *
            ( 0.5 * v[ijk - ii - jj] + v[ijk - ii] + v[ijk - ii + jj] + 0.5 * v[ijk - ii + jj2]
            + 0.5 * v[ijk      - jj] + v[ijk     ] + v[ijk      + jj] + 0.5 * v[ijk      + jj2]
            + 0.5 * v[ijk + ii - jj] + v[ijk + ii] + v[ijk + ii + jj] + 0.5 * v[ijk + ii + jj2] );
        const double du = u_filtered - 0.5 * (ubot[ij] + ubot[ij + ii]);
        const double dv = v_filtered - 0.5 * (vbot[ij] + vbot[ij + jj]);
        const double du2 = du * du + dv * dv;
        dutot[ij] = fmax(pow(du2, 0.5), 1.e-1);
    }

    // stability / stability_neutral, src/boundary_surface.cxx:54-176 — all ijcells, as the reference loops.
    // thermo = 0: neutral.  th-related pointers may be NULL then.
    __global__ void stability_kernel(double* __restrict__ ustar, double* __restrict__ obuk, int* __restrict__ nobuk,
                                     const double* __restrict__ th, const double* __restrict__ thbot,
                                     const double* __restrict__ thfluxbot, const double* __restrict__ dutot,
                                     const double* __restrict__ z0m, const double* __restrict__ z0h, const int constant_z0,
                                     const float* __restrict__ zL_sl,
                                     const float* __restrict__ f_sl, const int thermo, const int mbcbot, const int thermobc,
                                     const double g_thref, const double thref_k, const double g_threfh, const double threfh_k,
                                     const double db_ref, const double zsl, const GridD gd)
    {
        const int ij = blockIdx.x * blockDim.x + threadIdx.x;
        if (ij >= gd.ijcells) return;
        if (!thermo)
        {
            if (mbcbot == BC_USTAR)
            {
                const int i = ij % gd.icells, j = ij / gd.icells;     // interior only in this branch (:150-157)
                if (i >= gd.istart && i < gd.iend && j >= gd.jstart && j < gd.jend) obuk[ij] = -kDbig;
            }
            else if (mbcbot == BC_DIRICHLET)
            {
                obuk[ij] = -kDbig;
                ustar[ij] = dutot[ij] * most_fm(zsl, z0m[ij], -kDbig);
            }
            return;
        }
        if (mbcbot == BC_USTAR && thermobc == BC_FLUX)
        {
            const double us = ustar[ij];
            obuk[ij] = -(us * us * us) / (kKappa * (g_threfh * thfluxbot[ij]));
        }
        else if (mbcbot == BC_DIRICHLET && thermobc == BC_FLUX)
        {
            const double du = dutot[ij];
            double L;
            if (constant_z0)
            {
                const float Ri = -kKappa * (g_threfh * thfluxbot[ij]) * zsl / (du * du * du);
                int n = nobuk[ij];
                L = zsl / find_zL(zL_sl, f_sl, n, Ri);
                nobuk[ij] = n;
            }
            else
                L = obuk_iterative<true>(obuk[ij], du, g_threfh * thfluxbot[ij], zsl, z0m[ij], 0.);
            obuk[ij] = L;
            ustar[ij] = du * most_fm(zsl, z0m[ij], L);
        }
        else if (mbcbot == BC_DIRICHLET && thermobc == BC_DIRICHLET)
        {
            const double du = dutot[ij];
            const double b = g_thref * (th[ij + gd.kstart * gd.ijcells] - thref_k);      // calc_buoyancy_bot, src/thermo_dry.cxx:129-146
            const double bbot = g_threfh * (thbot[ij] - threfh_k);
            const double db = b - bbot + db_ref;
            double L;
            if (constant_z0)
            {
                const float Ri = kKappa * db * zsl / (du * du);
                int n = nobuk[ij];
                L = zsl / find_zL(zL_sl, f_sl, n, Ri);
                nobuk[ij] = n;
            }
            else
                L = obuk_iterative<false>(obuk[ij], du, db, zsl, z0m[ij], z0h[ij]);
            obuk[ij] = L;
            ustar[ij] = du * most_fm(zsl, z0m[ij], L);
        }
    }

    // surfm fluxes, src/boundary_surface.cxx:200-249 (interior; cyclic fill follows)
    __global__ void surfm_flux_kernel(double* __restrict__ ufluxbot, double* __restrict__ vfluxbot,
                                      const double* __restrict__ ustar, const double* __restrict__ obuk,
                                      const double* __restrict__ u, const double* __restrict__ ubot,
                                      const double* __restrict__ v, const double* __restrict__ vbot,
                                      const double* __restrict__ z0m, const double zsl, const int bcbot, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int ii = 1, jj = gd.icells;
        const int ij = i + j * jj, ijk = ij + gd.kstart * gd.ijcells;
        if (bcbot == BC_DIRICHLET)
        {
            const double c = ustar[ij] * most_fm(zsl, z0m[ij], obuk[ij]);
            ufluxbot[ij] = -(u[ijk] - ubot[ij]) * 0.5 * (ustar[ij - ii] * most_fm(zsl, z0m[ij - ii], obuk[ij - ii]) + c);
            vfluxbot[ij] = -(v[ijk] - vbot[ij]) * 0.5 * (ustar[ij - jj] * most_fm(zsl, z0m[ij - jj], obuk[ij - jj]) + c);
        }
        else if (bcbot == BC_USTAR)
        {
            const double minval = 1.e-2;
            auto p2 = [](double a) { return a * a; };
            auto p4 = [](double a) { return a * a * a * a; };
            const double vonu2 = fmax(minval, 0.25 * (p2(v[ijk - ii] - vbot[ij - ii]) + p2(v[ijk - ii + jj] - vbot[ij - ii + jj])
                                                    + p2(v[ijk] - vbot[ij]) + p2(v[ijk + jj] - vbot[ij + jj])));
            const double uonv2 = fmax(minval, 0.25 * (p2(u[ijk - jj] - ubot[ij - jj]) + p2(u[ijk + ii - jj] - ubot[ij + ii - jj])
                                                    + p2(u[ijk] - ubot[ij]) + p2(u[ijk + ii] - ubot[ij + ii])));
            const double u2 = fmax(minval, p2(u[ijk] - ubot[ij]));
            const double v2 = fmax(minval, p2(v[ijk] - vbot[ij]));
            const double ustaronu4 = 0.5 * (p4(ustar[ij - ii]) + p4(ustar[ij]));
            const double ustaronv4 = 0.5 * (p4(ustar[ij - jj]) + p4(ustar[ij]));
            ufluxbot[ij] = -copysign(1., u[ijk] - ubot[ij]) * pow(ustaronu4 / (1. + vonu2 / u2), 0.5);
            vfluxbot[ij] = -copysign(1., v[ijk] - vbot[ij]) * pow(ustaronv4 / (1. + uonv2 / v2), 0.5);
        }
    }

    // surfm gradients (:281-291) and surfs (:294-343): all ijcells
    __global__ void surf_grad_kernel(double* __restrict__ ugradbot, double* __restrict__ vgradbot, const double* __restrict__ u,
                                     const double* __restrict__ ubot, const double* __restrict__ v,
                                     const double* __restrict__ vbot, const double zsl, const GridD gd)
    {
        const int ij = blockIdx.x * blockDim.x + threadIdx.x;
        if (ij >= gd.ijcells) return;
        const int ijk = ij + gd.kstart * gd.ijcells;
        ugradbot[ij] = (u[ijk] - ubot[ij]) / zsl;
        vgradbot[ij] = (v[ijk] - vbot[ij]) / zsl;
    }

    __global__ void surfs_kernel(double* __restrict__ varbot, double* __restrict__ vargradbot, double* __restrict__ varfluxbot,
                                 const double* __restrict__ ustar, const double* __restrict__ obuk,
                                 const double* __restrict__ var, const double* __restrict__ z0h, const double zsl,
                                 const int bcbot, const GridD gd)
    {
        const int ij = blockIdx.x * blockDim.x + threadIdx.x;
        if (ij >= gd.ijcells) return;
        const int ijk = ij + gd.kstart * gd.ijcells;
        if (bcbot == BC_DIRICHLET)
        {
            varfluxbot[ij] = -(var[ijk] - varbot[ij]) * ustar[ij] * most_fh(zsl, z0h[ij], obuk[ij]);
            vargradbot[ij] = (var[ijk] - varbot[ij]) / zsl;
        }
        else if (bcbot == BC_FLUX)
        {
            varbot[ij] = varfluxbot[ij] / (ustar[ij] * most_fh(zsl, z0h[ij], obuk[ij])) + var[ijk];
            vargradbot[ij] = (var[ijk] - varbot[ij]) / zsl;
        }
    }

    // calc_duvdz_mo / calc_dbdz_mo, include/boundary_surface_kernels.h:190-247 (interior)
    __global__ void mo_grad_kernel(double* __restrict__ dudz, double* __restrict__ dvdz, double* __restrict__ dbdz,
                                   const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ ubot,
                                   const double* __restrict__ vbot, const double* __restrict__ thfluxbot,
                                   const double* __restrict__ ustar, const double* __restrict__ obuk,
                                   const double* __restrict__ z0m, const double g_threfh, const double zsl, const GridD gd)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x + gd.istart;
        const int j = blockIdx.y + gd.jstart;
        if (i >= gd.iend) return;
        const int ii = 1, jj = gd.icells;
        const int ij = i + j * jj, ijk = ij + gd.kstart * gd.ijcells;
        const double du_c = 0.5 * ((u[ijk] - ubot[ij]) + (u[ijk + ii] - ubot[ij + ii]));
        const double dv_c = 0.5 * ((v[ijk] - vbot[ij]) + (v[ijk + jj] - vbot[ij + jj]));
        const double fmv = most_fm(zsl, z0m[ij], obuk[ij]);
        const double ufl = -du_c * ustar[ij] * fmv;
        const double vfl = -dv_c * ustar[ij] * fmv;
        const double pm = phim(zsl / obuk[ij]);
        dudz[ij] = -ufl / (kKappa * zsl * ustar[ij]) * pm;
        dvdz[ij] = -vfl / (kKappa * zsl * ustar[ij]) * pm;
        if (dbdz)
            dbdz[ij] = -(g_threfh * thfluxbot[ij]) / (kKappa * zsl * ustar[ij]) * phih(zsl / obuk[ij]);
    }
}  // namespace

// Boundary_surface::init_solver -> prepare_lut, include/boundary_surface_kernels.h:78-138, on the host.
int surface_lut(Ctx* c, int mbcbot, int thermobc, double z0m, double z0h)
{
    const double zsl = c->z[c->gd.kstart];
    std::vector<double> zL_tmp(NLUT);
    std::vector<float> lut(2 * (size_t)NLUT, 0.f);
    float* zL_sl = lut.data();
    float* f_sl = lut.data() + NLUT;
    const double zLrange_min = -5.;
    double dzL = (kZLmax - zLrange_min) / (9. * NLUT / 10. - 1.);
    zL_tmp[0] = -kZLmax;
    for (int n = 1; n < 9 * NLUT / 10; ++n) zL_tmp[n] = zL_tmp[n - 1] + dzL;
    const double zLend = -(kZLmin - zLrange_min);
    double r = 1.01, r0 = kDhugeS;
    while (std::abs((r - r0) / r0) > 1.e-10)
    {
        r0 = r;
        r = std::pow(1. - (zLend / dzL) * (1. - r), (1. / (NLUT / 10.)));
    }
    for (int n = 9 * NLUT / 10; n < NLUT; ++n)
    {
        zL_tmp[n] = zL_tmp[n - 1] + dzL;
        dzL *= r;
    }
    for (int n = 0; n < NLUT; ++n) zL_sl[n] = -zL_tmp[NLUT - n - 1];
    if (mbcbot == BC_DIRICHLET && thermobc == BC_FLUX)
        for (int n = 0; n < NLUT; ++n) f_sl[n] = zL_sl[n] * std::pow(most_fm(zsl, z0m, zsl / zL_sl[n]), 3);
    else if (mbcbot == BC_DIRICHLET && thermobc == BC_DIRICHLET)
        for (int n = 0; n < NLUT; ++n)
            f_sl[n] = zL_sl[n] * std::pow(most_fm(zsl, z0m, zsl / zL_sl[n]), 2) / most_fh(zsl, z0h, zsl / zL_sl[n]);
    if (!c->surf_lut_d) MHH_CUDA_CHECK(cudaMalloc(&c->surf_lut_d, sizeof(float) * 2 * NLUT));
    MHH_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    MHH_CUDA_CHECK(cudaMemcpy(c->surf_lut_d, lut.data(), sizeof(float) * 2 * NLUT, cudaMemcpyHostToDevice));
    c->surf_lut_host = lut;
    return MHH_OK;
}

struct SurfaceArgs
{
    int mbcbot, thermo, thermobc;
    const double *u, *v;
    double *ubot, *vbot, *ufluxbot, *vfluxbot, *ugradbot, *vgradbot;
    const double* th;
    double *thbot, *thgradbot, *thfluxbot;
    double *ustar, *obuk;
    int* nobuk;
    const double *z0m, *z0h;
    double *dutot, *dudz, *dvdz, *dbdz;
    int constant_z0 = 1, charnock = 0;
    double alpha_m = 0., alpha_ch = 0., alpha_h = 0.;
};

// Boundary_surface<TF>::exec, src/boundary_surface.cxx:827-981 (swconstantz0 true: lookup solver, false: iterative solver
// on heterogeneous z0m / z0h, optionally updated by the Charnock relation)
// stages: 1 = total wind (+ its halo), 2 = stability + momentum fluxes (+ their halos), 4 = gradients / scalar / MO gradients.
// A y-slab rank (npy > 1) runs the stages one by one with the north/south exchange of dutot (after 1) and of the two flux
// planes (after 2) in between: the cyclic fills in here are east/west only on such a rank (cyclic_exec_k).
int surface_exec(Ctx* c, const SurfaceArgs& a, const int stages)
{
    const GridD& gd = c->gd;
    if (a.mbcbot == BC_NEUMANN) { c->last_error = "Neumann bc is not supported in surface model"; return MHH_ERR_ARG; }
    if (a.thermo && a.thermobc == BC_NEUMANN) { c->last_error = "Fixed Gradient bc is not supported in surface model"; return MHH_ERR_ARG; }
    if (a.thermo && a.thermobc == BC_DIRICHLET && a.mbcbot == BC_USTAR)
    { c->last_error = "Fixed Ustar bc in combination with Dirichlet bc for scalars is not supported"; return MHH_ERR_ARG; }
    if (a.thermo && !c->thref_d) { c->last_error = "thermo reference profiles not set"; return MHH_ERR_STATE; }
    if (a.thermo && a.mbcbot == BC_DIRICHLET && a.constant_z0 && !c->surf_lut_d) { c->last_error = "surface solver not prepared (mhh_surface_init_solver)"; return MHH_ERR_STATE; }
    if (a.charnock && a.constant_z0) { c->last_error = "\"swcharnock=true\" requires \"swconstantz0=false\""; return MHH_ERR_ARG; }
    if (!a.constant_z0 && !a.z0h) { c->last_error = "swconstantz0=false needs the z0h plane"; return MHH_ERR_ARG; }
    const double zsl = c->z[gd.kstart];
    const dim3 bi(128), gi((gd.imax + 127) / 128, gd.jmax);
    const int nb = (gd.ijcells + 255) / 256;
    int e;
    const double thref_k = a.thermo ? c->thref_ks : 1., threfh_k = a.thermo ? c->threfh_ks : 1.;
    const double g_thref = kGrav / thref_k, g_threfh = kGrav / threfh_k;
    const double db_ref = g_thref * (thref_k - threfh_k);          // Thermo_dry::get_db_ref, src/thermo_dry.cxx:629-633
    if ((stages & 1) && a.charnock)
    {
        // roughness lengths from the previous step's u* (src/boundary_surface.cxx:836-851); on a y-slab rank the caller
        // exchanges the north/south rows of z0m, z0h together with dutot's after this stage
        double* z0m = const_cast<double*>(a.z0m); double* z0h = const_cast<double*>(a.z0h);
        MHH_LAUNCH(c, "surf_charnock_kernel", charnock_kernel<<<gi, bi, 0, c->stream>>>(z0m, z0h, a.ustar, a.alpha_m, a.alpha_ch, a.alpha_h, gd));
        MHH_LAUNCH_CHECK();
        if ((e = cyclic_exec_2d(c, z0m))) return e;
        if ((e = cyclic_exec_2d(c, z0h))) return e;
    }
    if (stages & 1)
    {
        MHH_LAUNCH(c, "surf_dutot_kernel", dutot_kernel<<<gi, bi, 0, c->stream>>>(a.dutot, a.u, a.v, a.ubot, a.vbot, gd));
        MHH_LAUNCH_CHECK();
        if ((e = cyclic_exec_2d(c, a.dutot))) return e;
    }
    if (stages & 2)
    {
        MHH_LAUNCH(c, "surf_stability_kernel", stability_kernel<<<nb, 256, 0, c->stream>>>(
            a.ustar, a.obuk, a.nobuk, a.th, a.thbot, a.thfluxbot, a.dutot, a.z0m, a.z0h, a.constant_z0,
            c->surf_lut_d, c->surf_lut_d ? c->surf_lut_d + NLUT : nullptr,
            a.thermo, a.mbcbot, a.thermobc, g_thref, thref_k, g_threfh, threfh_k, db_ref, zsl, gd));
        MHH_LAUNCH_CHECK();
        MHH_LAUNCH(c, "surfm_flux_kernel", surfm_flux_kernel<<<gi, bi, 0, c->stream>>>(
            a.ufluxbot, a.vfluxbot, a.ustar, a.obuk, a.u, a.ubot, a.v, a.vbot, a.z0m, zsl, a.mbcbot, gd));
        MHH_LAUNCH_CHECK();
        if ((e = cyclic_exec_2d(c, a.ufluxbot))) return e;
        if ((e = cyclic_exec_2d(c, a.vfluxbot))) return e;
    }
    if (stages & 4)
    {
        MHH_LAUNCH(c, "surf_grad_kernel", surf_grad_kernel<<<nb, 256, 0, c->stream>>>(a.ugradbot, a.vgradbot, a.u, a.ubot, a.v, a.vbot, zsl, gd));
        MHH_LAUNCH_CHECK();
        if (a.thermo)
        {
            MHH_LAUNCH(c, "surfs_kernel", surfs_kernel<<<nb, 256, 0, c->stream>>>(
                a.thbot, a.thgradbot, a.thfluxbot, a.ustar, a.obuk, a.th, a.z0h, zsl, a.thermobc, gd));
            MHH_LAUNCH_CHECK();
        }
        MHH_LAUNCH(c, "surf_mo_grad_kernel", mo_grad_kernel<<<gi, bi, 0, c->stream>>>(
            a.dudz, a.dvdz, a.thermo ? a.dbdz : nullptr, a.u, a.v, a.ubot, a.vbot, a.thfluxbot, a.ustar, a.obuk, a.z0m, g_threfh, zsl, gd));
        MHH_LAUNCH_CHECK();
    }
    return MHH_OK;
}

int surface_scalar(Ctx* c, int bcbot, const double* var, double* varbot, double* vargradbot, double* varfluxbot,
                   const double* ustar, const double* obuk, const double* z0h)
{
    const GridD& gd = c->gd;
    if (bcbot == BC_NEUMANN) { c->last_error = "Fixed Gradient bc is not supported in surface model"; return MHH_ERR_ARG; }
    MHH_LAUNCH(c, "surfs_kernel", surfs_kernel<<<(gd.ijcells + 255) / 256, 256, 0, c->stream>>>(
        varbot, vargradbot, varfluxbot, ustar, obuk, var, z0h, c->z[gd.kstart], bcbot, gd));
    MHH_LAUNCH_CHECK();
    return MHH_OK;
}
}  // namespace mhh
