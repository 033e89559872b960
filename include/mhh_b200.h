/* mhh_b200.h — C ABI of the B200-native MicroHH dynamical core (libmhh_b200.so).
 *
 * The reference has no FFI; its CPU/GPU seam is compile-time substitution of the member
 * bodies guarded by `#ifndef USECUDA` (src/advec_2.cxx:265-336, src/diff_smag2.cxx:882-1189,
 * src/pres_2.cxx:64-105, src/timeloop.cxx:356-412, src/boundary.cxx:445-470,917-1008).
 * Each function below is what one of those guarded members binds to; INTEGRATION.md shows
 * the 3-line C++ shim per member.
 *
 * Conventions
 *  - All field pointers are DEVICE pointers to FP64 arrays in the reference layout:
 *    ijk = i + j*icells + k*icells*jcells, ghost cells included (src/grid.cxx:153-166).
 *    2-D planes are icells*jcells. Ownership stays with the caller (Field3d::fld_g,
 *    src/field3d.cu:31-61). Tendencies are accumulated into (+=), never overwritten.
 *  - Work is issued on the context stream (default: the legacy default stream, as the
 *    reference does); functions returning a host scalar synchronise that stream.
 *  - Return value: 0 on success, non-zero error code otherwise (the shim rethrows as
 *    std::runtime_error, matching include/tools.h:78-85). mhh_last_error() gives text.
 */
#ifndef MHH_B200_H
#define MHH_B200_H

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef struct mhh_ctx mhh_ctx;
/* always IEEE double, in the `_f32` symbol set too (include/mhh_b200_f32.h): time steps — the reference's rk3_g / rk4_g take
 * `mhh_double dt` for TF = float as well (src/timeloop.cu:35,74) and Advec/Diff/Pres take `mhh_double dt` (include/advec.h:57-58) */
#ifndef MHH_DOUBLE_DEFINED
#define MHH_DOUBLE_DEFINED
typedef double mhh_double;
#endif

/* Boundary_type, include/boundary.h:51 */
enum { MHH_BC_DIRICHLET = 0, MHH_BC_NEUMANN = 1, MHH_BC_FLUX = 2, MHH_BC_USTAR = 3, MHH_BC_OFF = 4 };
/* Boundary_w_type, include/boundary.h:52 */
enum { MHH_BCW_NORMAL = 0, MHH_BCW_CONSERVATION = 1 };
/* Edge, include/boundary_cyclic.h:33 */
enum { MHH_EDGE_EAST_WEST = 0, MHH_EDGE_NORTH_SOUTH = 1, MHH_EDGE_BOTH = 2 };
/* [advec] swadvec = 2 | 2i5 | 4 (src/advec.cxx:42-83) */
enum { MHH_ADVEC_2 = 2, MHH_ADVEC_2I5 = 25, MHH_ADVEC_4 = 4 };

/* Mirror of the scalars of Grid_data<TF> (include/grid.h:50-135) that define a grid.
 * npy/rank_y describe the y-slab decomposition (npx = 1): jmax = jtot/npy. */
typedef struct
{
    int itot, jtot, ktot;
    int igc, jgc, kgc;          /* ghost cells: 1,1,1 (order 2), 3,3,3 (order 4), 3,3,1|2 (advec 2i5) */
    double xsize, ysize, zsize;
    int order;                  /* [grid] swspatialorder: 2 or 4 */
    int npy, rank_y;            /* 1, 0 for a single GPU */
} mhh_grid_desc;

/* --- context / Grid (src/grid.cxx:106-189 init, :237-368 calculate) ------------------------ */
int mhh_create(const mhh_grid_desc* desc, const double* z_host /* ktot cell centres */, mhh_ctx** out);
int mhh_destroy(mhh_ctx* ctx);
const char* mhh_last_error(const mhh_ctx* ctx);
int mhh_set_stream(mhh_ctx* ctx, void* cuda_stream);
/* Fields::rhoref / rhorefh (kcells each, host). Default is 1 (Boussinesq, src/fields.cxx:394-395). */
int mhh_set_rhoref(mhh_ctx* ctx, const double* rhoref_host, const double* rhorefh_host);
/* out[20]: itot jtot ktot imax jmax kmax igc jgc kgc icells jcells kcells ijcells ncells
 *          istart iend jstart jend kstart kend */
int mhh_get_dims(const mhh_ctx* ctx, int* out);
/* name in {z,zh,dz,dzh,dzi,dzhi,dzi4,dzhi4,rhoref,rhorefh}: copies kcells doubles to host. */
int mhh_get_metric(const mhh_ctx* ctx, const char* name, double* out_host);
/* device pointer of the same arrays (for callers that launch their own kernels) */
const double* mhh_get_metric_device(const mhh_ctx* ctx, const char* name);

/* --- launch accounting and live per-kernel timing (no reference counterpart; used by bench.py) -------- */
unsigned long long mhh_launch_count(const mhh_ctx* ctx);          /* kernels launched so far by this ctx */
int mhh_profile_enable(mhh_ctx* ctx, int on);                     /* bracket every launch with CUDA events */
/* Sums recorded intervals per kernel since the last report. names: ';'-joined. Returns the entry count. */
int mhh_profile_report(mhh_ctx* ctx, char* names, int names_size, mhh_double* ms, long* counts, int max_entries);

/* --- Advec<TF> (include/advec.h:54-60) ------------------------------------------------------ */
/* Advec_2/2i5/4::exec — src/advec_2.cxx:298, src/advec_2i5.cxx:977, src/advec_4.cxx:592.
 * limit_flags[n] != 0 selects the flux-limited scalar scheme of include/advec_monotonic.h:77 (2i5 only). */
int mhh_advec_exec(mhh_ctx* ctx, int scheme,
                   const double* u, const double* v, const double* w,
                   double* ut, double* vt, double* wt,
                   int nscalars, const double* const* s, double* const* st, const int* limit_flags);
/* Advec_*::get_cfl — src/advec_2.cxx:267 (calc_cfl :51): cfl = dt * max(|u|/dx + |v|/dy + |w|/dz). */
int mhh_advec_cfl(mhh_ctx* ctx, int scheme, const double* u, const double* v, const double* w,
                  mhh_double dt, double* cfl_out);
/* Advec_*::get_time_limit — src/advec_2.cxx:281: idt * cflmax / max(cflmin, cfl). */
int mhh_advec_time_limit(mhh_ctx* ctx, int scheme, const double* u, const double* v, const double* w,
                         unsigned long idt, mhh_double dt, double cflmax, unsigned long* idtlim_out);

/* --- Diff<TF> (include/diff.h:45-64) ---------------------------------------------------------- */
/* Diff_2::exec src/diff_2.cxx:162 (diff_c :39 for u,v,scalars; diff_w :64). svisc[n] per scalar. */
int mhh_diff2_exec(mhh_ctx* ctx, const double* u, const double* v, const double* w,
                   double* ut, double* vt, double* wt, double visc,
                   int nscalars, const double* const* s, double* const* st, const double* svisc);
/* Diff_4::exec src/diff_4.cxx:253 (diff_c :42, diff_w :109). */
int mhh_diff4_exec(mhh_ctx* ctx, const double* u, const double* v, const double* w,
                   double* ut, double* vt, double* wt, double visc,
                   int nscalars, const double* const* s, double* const* st, const double* svisc);
/* Advec_4::exec, Boundary::set_ghost_cells_w(Normal_type) and Diff_4::exec (src/model.cxx:387-392) in one call: diff_c / diff_w
 * of u, v, w rides in the advection pass of the same component (no extra global traffic; the Normal-type w ghost values
 * diff_w needs are evaluated in the kernel), then the w ghost rows are switched and the scalars' diff_c follows.
 * w must hold Conservation-type ghost rows on entry (as for Advec::exec) and
 * holds Normal-type ones on return.  Bit-identical to the three separate calls. */
int mhh_advec4_diff4_exec(mhh_ctx* ctx, const double* u, const double* v, double* w, double* ut, double* vt, double* wt,
                          double visc, int nscalars, const double* const* s, double* const* st, const double* svisc);
/* Diff_2/4::get_dn: dnmul is static (src/diff_2.cxx:139-151). Returns dnmul (not * dt). */
int mhh_diff_const_dnmul(mhh_ctx* ctx, double viscmax, double* dnmul_out);

/* Diff_smag2::exec_viscosity — src/diff_smag2.cxx:1046 (calc_strain2 :48, calc_evisc :255,
 * calc_evisc_neutral :158), including the trailing boundary_cyclic.exec(evisc) and, for
 * resolved walls, the mirrored ghost rows.
 *  surface_model: boundary.get_switch() != "default";  N2 == NULL selects the neutral branch
 *  (thermo.get_switch() == "0").  dudz/dvdz/dbdz/z0m are 2-D planes (NULL when surface_model == 0). */
int mhh_smag2_exec_viscosity(mhh_ctx* ctx, int surface_model,
                             const double* u, const double* v, const double* w,
                             const double* N2, const double* dudz, const double* dvdz,
                             const double* dbdz, const double* z0m,
                             double cs, double tPr, double visc, double* evisc);
/* The same with Thermo_dry's N2 (calc_N2, src/thermo_dry.cxx:66-79; base state from mhh_thermo_set_ref) evaluated inside
 * the kernel from th, i.e. exec_viscosity's `thermo.get_thermo_field(*N2, "N2")` + calc_evisc (src/diff_smag2.cxx:1143-1160)
 * without materialising the N2 field. */
int mhh_smag2_exec_viscosity_dry(mhh_ctx* ctx, int surface_model,
                                 const double* u, const double* v, const double* w,
                                 const double* th, const double* dudz, const double* dvdz,
                                 const double* dbdz, const double* z0m,
                                 double cs, double tPr, double visc, double* evisc);
/* mhh_smag2_exec_viscosity_dry plus the two maxima Model::set_time_step (src/model.cxx:731-751) asks for right after
 * exec_viscosity of the first substep (src/model.cxx:354-357), taken in the same pass over u,v,w:
 *  cfl_over_dt = Advec_2::get_cfl's maximum before the `* dt` (calc_cfl, src/advec_2.cxx:51-78),
 *  dnmul       = Diff_smag2's calc_dnmul of the evisc just written (src/diff_smag2.cxx:712-735).
 * Equal to mhh_advec_cfl(scheme 2) / dt and mhh_smag2_dnmul run afterwards (same expressions, exact maxima). */
int mhh_smag2_exec_viscosity_dry_limits(mhh_ctx* ctx, int surface_model,
                                        const double* u, const double* v, const double* w,
                                        const double* th, const double* dudz, const double* dvdz,
                                        const double* dbdz, const double* z0m,
                                        double cs, double tPr, double visc, double* evisc,
                                        double* cfl_over_dt, double* dnmul);
/* mhh_advec2_smag2_exec_rk in parts: the momentum pass (do_uvw) and / or the scalar passes (do_scalars) restricted to the
 * 64-level k chunks [kc0, kc0 + nkc) (nkc <= 0: up to the top).  The scalar passes do not depend on the pressure solver; a
 * y-slab caller issues them on a second stream in halves that run while the copy engines carry the solver's all-to-all
 * chunks.  The parts of one substep together are bit-identical to one mhh_advec2_smag2_exec_rk call. */
int mhh_advec2_smag2_exec_rk_parts(mhh_ctx* ctx, int surface_model,
                                   const double* u, const double* v, const double* w, const double* evisc,
                                   double* ut, double* vt, double* wt,
                                   const double* ufluxbot, const double* ufluxtop, const double* vfluxbot, const double* vfluxtop,
                                   double visc, const double* th_buoy, int nscalars, const double* const* s, double* const* st,
                                   const double* const* sfluxbot, const double* const* sfluxtop, const double* svisc, double tPr,
                                   const double* const* sbuf, double* const* s_new, int rkorder, int substep, mhh_double dt,
                                   int do_uvw, int do_scalars, int kc0, int nkc);
/* Buffer::exec (src/buffer.cxx:163; calc_buffer :37-58) and the swlspres=geo branch of Force::exec (src/force.cxx:609-625,
 * calc_coriolis_2nd :76-107) for u,v,w, applied inside the fused momentum pass of mhh_advec2_smag2_exec[_rk] to the finished
 * tendency of each cell in the time loop's order (bit-identical to mhh_buffer_exec + mhh_force_coriolis run afterwards, without
 * their passes over u,v,ut,vt,wt).  ubuf/vbuf/wbuf: device profiles (kcells) or all NULL; sigma comes from mhh_buffer_plan.
 * ug/vg: device profiles (kcells).  The setting persists until changed; the caller then skips those two calls for u,v,w. */
int mhh_advec2_smag2_set_force_tail(mhh_ctx* ctx, const double* ubuf, const double* vbuf, const double* wbuf, int coriolis,
                                    double fc, const double* ug, const double* vg, double ugrid, double vgrid);
/* Tile-plane staging of the k-marching stencil kernels: TMA bulk tensor copies (cp.async.bulk.tensor, default where icells is
 * even) or per-thread cp.async (on = 0).  Both give identical results; tests compare them. */
int mhh_set_tma(mhh_ctx* ctx, int on);
int mhh_tma_active(mhh_ctx* ctx);
/* Lazy-tendency mode of the fused 2nd-order LES path (off by default).  Timeloop::exec leaves every tendency at zero after
 * the last substep (`at *= cA`, cA = 0: src/timeloop.cxx:250-285) and the next step's first tendency operator reads the zeros
 * back.  With the mode on, the last-substep tails (mhh_pres_exec_rk: ut,vt,wt; mhh_advec2_smag2_exec_rk: the scalars' st)
 * leave those arrays UNDEFINED, and the next mhh_advec2_smag2_exec[_rk] / mhh_smag2_exec call on the same arrays starts from 0
 * without reading them.  Results are bit-identical; nothing else may read those arrays between the two calls. */
int mhh_set_lazy_tendencies(mhh_ctx* ctx, int on);
/* mhh_pres_exec / mhh_pres_exec_rk (2nd order, power-of-two itot, even jmax): evaluate Pres_2::input (src/pres_2.cxx:183-195)
 * inside the forward x transform instead of writing the right-hand side and reading it back (bit-identical; off by default). */
int mhh_pres_set_fused_input(mhh_ctx* ctx, int on);
/* Diff_smag2::exec — src/diff_smag2.cxx:939 (diff_u :370, diff_v :472, diff_w :575, diff_c :620). */
int mhh_smag2_exec(mhh_ctx* ctx, int surface_model,
                   const double* u, const double* v, const double* w, const double* evisc,
                   double* ut, double* vt, double* wt,
                   const double* ufluxbot, const double* ufluxtop,
                   const double* vfluxbot, const double* vfluxtop, double visc,
                   int nscalars, const double* const* s, double* const* st,
                   const double* const* sfluxbot, const double* const* sfluxtop,
                   const double* svisc, double tPr);
/* One pass per prognostic field for the LES tendencies (no reference counterpart): Thermo_dry::exec's buoyancy on wt
 * (src/thermo_dry.cxx:43-62; only when th_buoy != NULL), Advec_2::exec (src/advec_2.cxx:98-202) and Diff_smag2::exec, summed
 * in the time loop's order so that the result is bit-identical to mhh_thermo_dry_exec, mhh_advec_exec, mhh_smag2_exec. */
int mhh_advec2_smag2_exec(mhh_ctx* ctx, int surface_model,
                          const double* u, const double* v, const double* w, const double* evisc,
                          double* ut, double* vt, double* wt,
                          const double* ufluxbot, const double* ufluxtop,
                          const double* vfluxbot, const double* vfluxtop, double visc,
                          const double* th_buoy,
                          int nscalars, const double* const* s, double* const* st,
                          const double* const* sfluxbot, const double* const* sfluxtop,
                          const double* svisc, double tPr);
/* The same pass with, per scalar, the two operators that follow Diff::exec in Model::exec when nothing else adds to the
 * scalar's tendency (no [force] term on it), applied to the finished tendency before it is stored:
 *   sbuf[n] != NULL : Buffer::exec of that scalar (src/buffer.cxx:37-58; needs mhh_buffer_plan), profile sbuf[n];
 *   s_new[n] != NULL: Timeloop::exec of that scalar (src/timeloop.cxx:250-304): s_new = s + cB dt st, st = cA st (zeroed,
 *                     ghost cells included, after the last substep).  s_new[n] is a second array of the same size — the
 *                     interior holds the updated field on return, the caller swaps the two pointers and refreshes the
 *                     ghost cells as at the start of every substep (src/model.cxx:346-348).
 * sbuf / s_new may be NULL (then identical to mhh_advec2_smag2_exec).  Bit-identical to the separate calls; saves 3 of the 4
 * field passes of the scalar's RK update (and the 3 of its Buffer pass). */
int mhh_advec2_smag2_exec_rk(mhh_ctx* ctx, int surface_model,
                             const double* u, const double* v, const double* w, const double* evisc,
                             double* ut, double* vt, double* wt,
                             const double* ufluxbot, const double* ufluxtop,
                             const double* vfluxbot, const double* vfluxtop, double visc,
                             const double* th_buoy,
                             int nscalars, const double* const* s, double* const* st,
                             const double* const* sfluxbot, const double* const* sfluxtop,
                             const double* svisc, double tPr,
                             const double* const* sbuf, double* const* s_new, int rkorder, int substep, mhh_double dt);
/* calc_strain2 alone (src/diff_smag2.cxx:48) — stage check for parity tests; the product path fuses it. */
int mhh_smag2_strain2(mhh_ctx* ctx, int surface_model, const double* u, const double* v, const double* w,
                      const double* dudz, const double* dvdz, double* strain2);
/* calc_dnmul — src/diff_smag2.cxx:712; get_dn :904 multiplies by dt, get_time_limit :884. */
int mhh_smag2_dnmul(mhh_ctx* ctx, const double* evisc, double tPr, double* dnmul_out);

/* --- Thermo_dry pieces inside the substep ---------------------------------------- */
/* base state thref/threfh (kcells each, host): Thermo_dry::create_basestate, src/thermo_dry.cxx */
int mhh_thermo_set_ref(mhh_ctx* ctx, const double* thref_host, const double* threfh_host);
/* Thermo_dry::exec — calc_buoyancy_tend_2nd/4th, src/thermo_dry.cxx:165-198,459-491: wt += g/threfh*(th_h - threfh) */
int mhh_thermo_dry_exec(mhh_ctx* ctx, double* wt, const double* th);
/* Thermo_dry::get_thermo_field("N2") — calc_N2, src/thermo_dry.cxx:66-79 */
int mhh_thermo_dry_N2(mhh_ctx* ctx, const double* th, double* N2);

/* --- Boundary_surface<TF> (Monin-Obukhov surface layer; include/boundary_surface.h, src/boundary_surface.cxx) —
 * SURVEY 8(f) rank 4.  swconstantz0=true (lookup solver), no Charnock; bc codes as MHH_BC_*. --------------------------- */
/* Boundary_surface::init_solver -> prepare_lut (include/boundary_surface_kernels.h:78-138), built on the host */
int mhh_surface_init_solver(mhh_ctx* ctx, int mbcbot, int thermobc, double z0m, double z0h);
/* copies the table out (2 x 10000 floats: zL_sl, f_sl) — for tests and restart checks */
int mhh_surface_get_lut(mhh_ctx* ctx, float* zL_sl_host, float* f_sl_host);
typedef struct mhh_surface_args
{
    int mbcbot;              /* MHH_BC_DIRICHLET (noslip) or MHH_BC_USTAR */
    int thermo;              /* 0: thermo.get_switch() == "0" (stability_neutral); 1: dry (needs mhh_thermo_set_ref) */
    int thermobc;            /* bottom bc of th: MHH_BC_DIRICHLET or MHH_BC_FLUX */
    const double *u, *v;     /* 3-D */
    double *ubot, *vbot, *ufluxbot, *vfluxbot, *ugradbot, *vgradbot;      /* Field3d planes of u, v */
    const double* th;        /* 3-D (NULL when thermo == 0) */
    double *thbot, *thgradbot, *thfluxbot;                                /* Field3d planes of th */
    double *ustar, *obuk;    /* Boundary_surface::ustar, obuk (ijcells; persist between calls) */
    int* nobuk;              /* Boundary_surface::nobuk (ijcells ints; the lookup resumes where it ended) */
    const double *z0m, *z0h; /* ijcells */
    double* dutot;           /* ijcells scratch (the reference borrows a tmp field) */
    double *dudz, *dvdz, *dbdz;   /* dudz_mo, dvdz_mo, dbdz_mo: the bottom-row inputs of mhh_smag2_exec_viscosity */
    /* [boundary] swconstantz0=false: the iterative Obukhov-length solver on heterogeneous z0m / z0h planes
     * (calc_obuk_noslip_*_iterative, include/boundary_surface_kernels.h:287-470) instead of the z/L lookup; 0 = lookup.
     * swcharnock=true (needs the iterative solver): z0m, z0h are first updated from the previous u* (calc_z0_charnock,
     * src/boundary_surface.cxx:342-365) with the three [boundary] alpha_* constants, and must then be writable. */
    int iterative_solver, charnock;
    double alpha_m, alpha_ch, alpha_h;
} mhh_surface_args;
/* Boundary_surface::exec, src/boundary_surface.cxx:827-981: calc_dutot, stability / stability_neutral, surfm, surfs(th),
 * calc_duvdz_mo, calc_dbdz_mo with the cyclic fills in between. */
int mhh_boundary_surface_exec(mhh_ctx* ctx, const mhh_surface_args* args);
/* The same in stages for a y-slab rank (the MPI build's boundary_cyclic.exec_2d calls inside Boundary_surface::exec,
 * src/boundary_surface.cxx:850,930-931, become neighbour exchanges between the stages): stages is a mask of
 * 1 = calc_dutot (+ east/west fill), 2 = stability + surfm fluxes (+ east/west fills of ufluxbot, vfluxbot), 4 = the rest.
 * The caller exchanges the north/south rows of dutot after stage 1 and of ufluxbot, vfluxbot after stage 2. */
int mhh_boundary_surface_exec_stages(mhh_ctx* ctx, const mhh_surface_args* args, int stages);
/* surfs for a further scalar, src/boundary_surface.cxx:294-343,935-946 */
int mhh_surface_scalar(mhh_ctx* ctx, int bcbot, const double* var, double* varbot, double* vargradbot, double* varfluxbot,
                       const double* ustar, const double* obuk, const double* z0h);

/* --- Buffer<TF> (sponge layer; include/buffer.h, src/buffer.cxx) — SURVEY 8(f) rank 3 --------------------- */
/* Buffer::create, src/buffer.cxx:101-125: finds bufferkstart/bufferkstarth (first full/half level with z >= zstart),
 * fails with "Buffer is too close to the model top" like the reference, and tabulates sigma*((z-zstart)/(zsize-zstart))^beta
 * per level on the host (std::pow, :49). */
int mhh_buffer_plan(mhh_ctx* ctx, double zstart, double sigma, double beta, int* bufferkstart, int* bufferkstarth);
/* calc_buffer, src/buffer.cxx:37-58 (buffer_g, src/buffer.cu:36-60): at -= sigma(z)*(a - abuf[k]) above zstart.
 * abuf: device profile (kcells) — bufferprofs[name] or fld_mean when swupdate; is_w selects the half-level table. */
int mhh_buffer_exec(mhh_ctx* ctx, double* at, const double* a, const double* abuf, int is_w);

/* --- Force<TF>::exec (include/force.h, src/force.cxx:586-720) ---------------------------------------------- */
/* swlspres=dpdx: add_pressure_force(ut, -dpdx), src/force.cxx:46-60,603-611 */
int mhh_force_pressure(mhh_ctx* ctx, double* ut, double fbody);
/* swlspres=uflux: enforce_fixed_flux with the two calc_mean reductions, src/force.cxx:63-75,590-601 */
int mhh_force_fixed_flux(mhh_ctx* ctx, double* ut, const double* u, double uflux, double utrans, mhh_double dt, double* fbody_out);
/* swlspres=geo: calc_coriolis_2nd/_4th, src/force.cxx:78-144,613-634; ug, vg device profiles (kcells) */
int mhh_force_coriolis(mhh_ctx* ctx, double* ut, double* vt, const double* u, const double* v,
                       const double* ug, const double* vg, double fc, double utrans, double vtrans);
/* swls=1: calc_large_scale_source, src/force.cxx:146-161 (sls: device profile) */
int mhh_force_large_scale_source(mhh_ctx* ctx, double* st, const double* sls);
/* swnudge=1: calc_nudging_tendency, src/force.cxx:163-183 (fldmean, ref, factor: device profiles) */
int mhh_force_nudge(mhh_ctx* ctx, double* at, const double* fldmean, const double* ref, const double* factor);
/* swwls: mode 0 = advec_wls_2nd_mean (s = mean profile), 1 = advec_wls_2nd_local, 2 = advec_wls_2nd_local_w
 * (s = 3-D field), src/force.cxx:195-292 */
int mhh_force_subsidence(mhh_ctx* ctx, int mode, double* st, const double* s, const double* wls);

/* --- Pres<TF> (include/pres.h:48-58) --------------------------------------------------------------- */
/* Pres_2/4::init + set_values — src/pres_2.cxx:107-153, src/pres_4.cxx:159-252. Uses ctx rhoref. */
int mhh_pres_set_values(mhh_ctx* ctx);
/* Pres_2/4::exec(dt) — src/pres_2.cxx:66, src/pres_4.cxx:77: input, solve, output. p is ghosted. */
int mhh_pres_exec(mhh_ctx* ctx, const double* u, const double* v, const double* w,
                  double* ut, double* vt, double* wt, double* p, mhh_double dt);
/* Pres_2/4::check_divergence — src/pres_2.cxx:99,391; src/pres_4.cxx:144,733. */
int mhh_pres_check_divergence(mhh_ctx* ctx, const double* u, const double* v, const double* w,
                              double* divmax_out);
/* Stages, exposed for parity tests and for the multi-GPU driver (transposes happen between them):
 *  input : Pres_2::input src/pres_2.cxx:156 -> ghost-free work array (device, imax*jmax*kmax)
 *  fft   : FFT::exec_forward/backward src/fft.cxx:588-599, half-complex in place
 *  solve : solve_in + tdma/hdma src/pres_2.cxx:292-327, src/pres_4.cxx:320-505
 *  output: scatter + ghost rows + cyclic + Pres_*::output src/pres_2.cxx:340-387 */
double* mhh_pres_work(mhh_ctx* ctx);
int mhh_pres_input(mhh_ctx* ctx, const double* u, const double* v, const double* w,
                   double* ut, double* vt, double* wt, mhh_double dt);
int mhh_fft_force_generic(mhh_ctx* ctx, int on);   /* tests: use the mixed-radix kernel even for 2^k sizes */
int mhh_fft_forward(mhh_ctx* ctx, double* data);
int mhh_fft_backward(mhh_ctx* ctx, double* data);      /* normalised by 1/(itot*jtot), in place */
int mhh_pres_solve_spectral(mhh_ctx* ctx, double* data);
int mhh_pres_output(mhh_ctx* ctx, double* ut, double* vt, double* wt, double* p);
/* y-slab (multi-GPU) stages. Layouts: X = [k][j<jmax][i<itot] (slab), Y = [k][j<jtot][ii<iblock];
 * transpose modes: 0 X->chunks, 1 chunks->Y (Transpose::exec_xy, src/transpose.cxx:163-190),
 *                  2 Y->chunks, 3 chunks->X (Transpose::exec_yx, :191-217); chunk r = [k][j<jmax][ii<iblock].
 * The all-to-all between pack and unpack is the host's (NCCL over NVLink). dir: -1 forward, +1 backward. */
int mhh_pres_input_to(mhh_ctx* ctx, const double* u, const double* v, const double* w, const double* ut,
                      const double* vt, const double* wt, mhh_double dt, double* dst_X);
int mhh_fft_x(mhh_ctx* ctx, int dir, double* data_X, double* p_or_null);
int mhh_fft_y(mhh_ctx* ctx, int dir, double* data_Y);
/* FFT stages that read / write the all-to-all chunk layout directly (chunk for peer r = [k][j<jmax][ii<iblock]), fusing
 * Transpose::exec_xy / exec_yx's pack and unpack copies (src/transpose.cxx:163-217) into the transforms:
 *   x, dir<0: slab X -> chunks;  y, dir<0: received chunks -> Y;  y, dir>0: Y -> chunks;  x, dir>0: received chunks -> X or p
 * [k0, k0+nk) (nk = 0: all levels) selects a k range whose chunk buffer is laid out on its own,
 * [k0..][r][k'<nk][j<jmax][ii<iblock], so that the collective of one range overlaps the transforms of the next. */
int mhh_fft_x_chunks(mhh_ctx* ctx, int dir, const double* in, double* out, double* p_or_null, int k0, int nk);
int mhh_fft_y_chunks(mhh_ctx* ctx, int dir, const double* in, double* out, int k0, int nk);
int mhh_fft_chunks_supported(mhh_ctx* ctx);

/* --- peer-memory data path of the y-slab decomposition (no reference counterpart: replaces the MPI messages of
 * Boundary_cyclic::exec, src/boundary_cyclic.cxx:139-174, and of Transpose::exec_xy/exec_yx, src/transpose.cxx:163-217) ---
 * Every rank allocates one exchange arena (mhh_peer_alloc: cudaMalloc, zeroed), places its halo-exchanged fields, the solver's
 * receive buffers and 32 flag words in it at the SAME offsets on all ranks, exports it (64-byte CUDA IPC handle), imports the
 * other ranks' handles and registers the table of bases.  From then on data moves with stores to peer addresses: */
int mhh_peer_alloc(mhh_ctx* ctx, unsigned long long bytes, void** ptr);
int mhh_peer_free(mhh_ctx* ctx, void* ptr);
int mhh_peer_export(mhh_ctx* ctx, void* ptr, unsigned char* handle64);
int mhh_peer_import(mhh_ctx* ctx, const unsigned char* handle64, void** mapped_ptr);
int mhh_peer_unmap(mhh_ctx* ctx, void* mapped_ptr);
/* bases[r]: rank r's arena as mapped in this process (bases[me] == my_base); flags_offset: byte offset of the 32 flag words. */
int mhh_peer_register(mhh_ctx* ctx, int world, int me, void* my_base, unsigned long long bytes, void* const* bases,
                      unsigned long long flags_offset);
/* North/south halo rows of up to 8 fields (3-D, or 2-D planes with is2d): my top interior rows are stored into the north
 * neighbour's south ghost rows and my bottom rows into the south neighbour's north ghost rows; returns (stream-ordered)
 * when both neighbours' rows have arrived in mine.  The east/west fill (mhh_cyclic) comes first, so corners are right. */
int mhh_peer_halo_push(mhh_ctx* ctx, int nfields, double* const* flds, int is2d);
/* All ranks have completed and made visible everything stream-ordered before this call. */
int mhh_peer_barrier(mhh_ctx* ctx);
int mhh_peer_error(mhh_ctx* ctx, int* err);    /* 1 after a flag wait timed out (a peer was lost) */
/* Forward x transform / inverse y transform whose chunk-layout output (mhh_fft_x_chunks dir < 0, mhh_fft_y_chunks dir > 0) is
 * written straight into the peers' receive buffers — peer_out[r] = rank r's buffer as mapped here; segment r of every line
 * lands in slot `rank_y` of rank r's buffer: the all-to-all happens inside the transform, tile by tile over NVLink.  Follow
 * with mhh_peer_barrier before the receiving transform (mhh_fft_y_chunks dir < 0 / mhh_fft_x_chunks dir > 0) reads. */
int mhh_fft_x_chunks_to_peers(mhh_ctx* ctx, const double* in, double* const* peer_out, int k0, int nk);
int mhh_fft_y_chunks_to_peers(mhh_ctx* ctx, const double* in, double* const* peer_out, int k0, int nk);
/* The solver's transforms on kx SUB-BLOCKS of a rank's block (nsub per rank, ib = iblock / nsub columns each; nsub divides
 * iblock, ranks * nsub <= 32), so that the y transforms and the vertical solve of sub-block b run while the copy engines move
 * sub-blocks b+1 (forward) and b-1 (return) — Pres_2::solve (src/pres_2.cxx:267-349) stops being a join between the two
 * transposes (src/transpose.cxx:163-217).  Segment v = r * nsub + b of an x line (rank r's sub-block b) is
 * [k][j < jmax][ii < ib], Rb = kmax * jmax * ib elements; a receive buffer holds segment v (v over the source ranks) at v * Rb;
 * Y_b = [k][j < jtot][ii < ib] sits at b * kmax * jtot * ib of the Y array.
 *   x_fwd : X -> segments, segment v of every line written at segptrs[v] (v < ranks * nsub: staging array or a buffer itself)
 *   y_fwd : the received segments of sub-block b -> Y_b            solve : Thomas sweep of Y_b (Pres_2; work >= Y_b / 16 + a plane)
 *   y_bwd : Y_b -> segments by destination rank r at segptrs[r]     x_bwd : all received segments -> the ghost-celled p */
int mhh_slab_fft_x_fwd_sub(mhh_ctx* ctx, const double* in_X, double* const* segptrs, int nsub);
int mhh_slab_fft_y_fwd_sub(mhh_ctx* ctx, const double* in_segments, double* out_Y, int b, int nsub);
int mhh_slab_solve_sub(mhh_ctx* ctx, double* Y, double* work, int b, int nsub);
int mhh_slab_fft_y_bwd_sub(mhh_ctx* ctx, const double* in_Y, double* const* segptrs, int b, int nsub);
int mhh_slab_fft_x_bwd_sub(mhh_ctx* ctx, const double* in_segments, double* p, int nsub);
int mhh_transpose_copy(mhh_ctx* ctx, int mode, double* dst, const double* src);
int mhh_pres_solve_spectral_ws(mhh_ctx* ctx, double* data_Y, double* work3d);
int mhh_pres_output_stage(mhh_ctx* ctx, int stage, double* ut, double* vt, double* wt, double* p);
/* north/south halo rows of a ghosted field <-> [k][g<jgc][i<icells] buffers (Boundary_cyclic MPI path,
 * src/boundary_cyclic.cxx:139-174). unpack=0: north<-rows [jend-jgc,jend), south<-rows [jstart,jstart+jgc);
 * unpack=1: rows [jend,jend+jgc)<-north buffer, rows [0,jgc)<-south buffer. */
int mhh_halo_ns(mhh_ctx* ctx, double* fld, double* north_buf, double* south_buf, int unpack, int is2d);
int mhh_pres_scatter(mhh_ctx* ctx, double* p, const double* ghost_free_src);   /* src/pres_2.cxx:340-348 */
/* copies bmati(itot), bmatj(jtot) and a,c (kmax each; order 2) or m1..m7 (7*kmax; order 4) to host */
int mhh_pres_get_coefs(const mhh_ctx* ctx, double* bmati, double* bmatj, double* a_or_bands, double* c);

/* --- Boundary<TF> / Boundary_cyclic<TF> (include/boundary.h:92-98, include/boundary_cyclic.h) ---------- */
/* Boundary_cyclic::exec / exec_2d — src/boundary_cyclic.cxx:369,445. */
int mhh_cyclic(mhh_ctx* ctx, double* fld, int edge);
int mhh_cyclic_2d(mhh_ctx* ctx, double* fld2d);
/* calc_ghost_cells_{bot,top}_{2nd,4th} — src/boundary.cxx:687-836, one field; order from ctx. */
int mhh_ghost_cells(mhh_ctx* ctx, double* a, int bcbot, int bctop,
                    double* abot, double* atop, const double* agradbot, const double* agradtop);
/* Boundary::set_ghost_cells_w — src/boundary.cxx:986 (no-op for order 2). */
int mhh_ghost_cells_w(mhh_ctx* ctx, double* w, int type);
/* Boundary::set_values for one field: fill the 2-D planes from scalar BC values
 * (set_bc, src/boundary.cxx:51-91).  visc and no_offset as in the reference. */
int mhh_set_bc(mhh_ctx* ctx, double* a2d, double* agrad2d, double* aflux2d, int sw, double aval,
               double visc, double offset);

/* --- Timeloop<TF>::exec (src/timeloop.cxx:358; rk3 :250, rk4 :288) -------------------------------------- */
int mhh_rk_substep(mhh_ctx* ctx, int rkorder, int substep, mhh_double dt,
                   int nfields, double* const* a, double* const* at);
/* Fused tail of the substep (no reference counterpart: Pres_2::output src/pres_2.cxx:365-387 / Pres_4::output
 * src/pres_4.cxx:533-571 immediately followed by Timeloop::exec src/timeloop.cxx:358-411 for u,v,w).  Bit-identical to
 * mhh_pres_output / mhh_pres_exec followed by mhh_rk_substep on (u,v,w); the caller runs mhh_rk_substep on the remaining
 * fields with the same substep.  4th order (mhh_pres_exec_rk only): the call also performs
 * Boundary::set_ghost_cells_w(Normal) on w (src/model.cxx:412) before the update, so the caller must not repeat it. */
int mhh_pres_output_rk(mhh_ctx* ctx, double* ut, double* vt, double* wt, double* p, double* u, double* v, double* w,
                       int rkorder, int substep, mhh_double dt);
int mhh_pres_exec_rk(mhh_ctx* ctx, double* u, double* v, double* w, double* ut, double* vt, double* wt, double* p,
                     mhh_double dt_sub, int rkorder, int substep, mhh_double dt);
/* One-shot: the next mhh_pres_exec_rk makes its stream wait for `cuda_event` (a cudaEvent_t) right before the kernel that
 * updates u, v, w — so that independent work issued on ANOTHER stream that still reads the old velocities (the scalars'
 * tendency + RK pass, which Model::exec orders before Pres::exec, src/model.cxx:388-411) may overlap the solver's input,
 * transforms and vertical solve.  The caller records the event on that other stream after the work. */
int mhh_pres_set_output_wait(mhh_ctx* ctx, void* cuda_event);
mhh_double mhh_rk_sub_dt(int rkorder, int substep, mhh_double dt);   /* Timeloop::get_sub_time_step :415 */

/* --- Field3d_operators<TF> (include/field3d_operators.h:40-51) and Fields diagnostics -------------------- */
int mhh_mean_profile(mhh_ctx* ctx, const double* fld, double* prof_host /* kcells */);
int mhh_level_sums(mhh_ctx* ctx, const double* fld, double* sums_host /* kcells, this rank only */);
/* Stats profiles over the default mask (every point), the observables of "stats profiles after N steps": src/stats.cxx
 * calc_mean :264, calc_moment :342 (sum of (fld - fld_mean[k] + offset)^power), calc_cov :368 (the product of two such
 * factors, e.g. w'th').  Host profiles of kcells values; levels kstart..kend are filled (half-level fields own level kend),
 * each divided by the GLOBAL itot*jtot: on a y-slab rank the profiles of the ranks add up (master.sum, src/stats.cxx:1887). */
int mhh_stats_mean(mhh_ctx* ctx, const double* fld, double* prof_host);
int mhh_stats_moment(mhh_ctx* ctx, const double* fld, const double* fld_mean_host, double offset, int power, double* prof_host);
int mhh_stats_cov(mhh_ctx* ctx, const double* fld1, const double* mean1_host, double offset1, int pow1,
                  const double* fld2, const double* mean2_host, double offset2, int pow2, double* prof_host);
int mhh_mean(mhh_ctx* ctx, const double* fld, double* mean_out);
int mhh_mean_2d(mhh_ctx* ctx, const double* fld2d, double* mean_out);
int mhh_max(mhh_ctx* ctx, const double* fld, double* max_out);   /* calc_max_g: max over interior */
/* No counterpart in the reference.  mhh_mean for callers that must not
 * allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the per-level partial sums live in a
 * file-scope static __device__ array of 4096 FP64 slots (32 KiB per GPU that loads the library, held until exit) instead of
 * the context's cudaMalloc'ed scratch.  fld and mean_out are DEVICE pointers (mean_out: one value); the two kernels are
 * queued on the context's stream, nothing is waited for, and no cudaMalloc / cudaFree / allocating routine is called.
 * ktot > 4096 is rejected with MHH_ERR_ARG (mhh_last_error says so).  All calls on one device share the array: calls that
 * overlap on the GPU race; consecutive calls on one stream do not (every slot is rewritten before it is read). */
int mhh_mean_static(mhh_ctx* ctx, const double* fld_dev, double* mean_out_dev);
/* Fields::check_momentum/tke/mass — src/fields.cxx:144,175,209 */
int mhh_check_momentum(mhh_ctx* ctx, const double* u, const double* v, const double* w, double* out);
int mhh_check_tke(mhh_ctx* ctx, const double* u, const double* v, const double* w, double* out);
int mhh_check_mass(mhh_ctx* ctx, const double* s, double* out);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
