/* topoflow_b200.h -- C ABI of libtopoflow_b200.so
 *
 * B200 (sm_100a) implementation of TopoFlow 3.5's per-timestep D8 channel
 * routing and of the D8 grid builder it runs on.  The reference
 * (peckhams/topoflow_old) is pure Python/NumPy and has no FFI of its own; each
 * entry point below names the reference method(s) it stands in for, as
 * topoflow/<file>:<line>.  The Python shim in topoflow_old_b200/ binds these
 * with ctypes; INTEGRATION.md shows the same binding written inside the
 * reference's own components.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; the message
 *     is available from tf_last_error() (thread-local to the calling thread);
 *   - all pointers are HOST pointers to dense row-major [ny][nx] grids (row 0
 *     = north) unless a parameter says otherwise; sizes are in elements;
 *   - a handle owns every device buffer; nothing is shared between handles;
 *   - one handle = one row strip of a (possibly larger) global grid on one
 *     GPU.  With a single strip the strip is the grid.
 */
#ifndef TOPOFLOW_B200_H
#define TOPOFLOW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tf_handle_s* tf_handle;

/* ---- selectors ---------------------------------------------------------- */
enum tf_wave { TF_KINEMATIC = 0, TF_DIFFUSIVE = 1, TF_DYNAMIC = 2 };
enum tf_law  { TF_MANNING = 0, TF_LAW_OF_WALL = 1 };

/* static channel parameters (channels_base.py:2142-2200, :797-984) */
enum tf_param {
    TF_P_WIDTH = 0,     /* bottom width [m], after "width == 0 -> dw"        */
    TF_P_TAN_ANGLE = 1, /* tan(bank angle), host libm (np.tan)               */
    TF_P_COS_ANGLE = 2, /* cos(bank angle), host libm (np.cos)               */
    TF_P_SLOPE = 3,     /* S_bed = slope / sinu, after remove_bad_slopes     */
    TF_P_ROUGH = 4,     /* nval (Manning) or z0val (law of the wall)         */
    TF_P_DS = 5,        /* sinu * d8.ds, the channel flow length [m]         */
    TF_P_COUNT = 6
};

/* grids that tf_chan_get / tf_chan_set_state understand */
enum tf_field {
    TF_F_VOL = 0, TF_F_D = 1, TF_F_U = 2, TF_F_Q = 3,   /* materialised state */
    TF_F_QNEXT = 4,     /* u * A_wet of the last step = next step's Q         */
    TF_F_F = 5, TF_F_RH = 6, TF_F_A_WET = 7, TF_F_P_WET = 8, TF_F_TAU = 9,
    TF_F_U_STAR = 10, TF_F_FROUDE = 11, TF_F_S_FREE = 12,   /* derived on call */
    TF_F_WIDTH = 13,    /* width with the overbank substitution applied       */
    TF_F_MUTATED = 14,  /* 1.0 where "top_width > dw" has fired, else 0.0     */
    TF_F_R = 15,        /* excess rain rate grid (only while R is a grid)      */
    TF_F_COUNT = 16
};

/* the six fluxes update_R combines (channels_base.py:1103-1147), in the
 * order the reference adds them: R = (P_rain + SM + GW + MR) - (ET + IN) */
enum tf_forcing { TF_R_P_RAIN = 0, TF_R_SM = 1, TF_R_GW = 2, TF_R_MR = 3,
                  TF_R_ET = 4, TF_R_IN = 5, TF_R_COUNT = 6 };

/* D8 products that tf_d8_get understands (element type in brackets) */
enum tf_d8_field {
    TF_D8_CODE = 0,       /* [uint8]   d8_grid                               */
    TF_D8_PARENT_ID = 1,  /* [int32]   parent_ID_grid                        */
    TF_D8_DS = 2,         /* [float32] flow length                           */
    TF_D8_DW = 3,         /* [float32] flow width                            */
    TF_D8_AREA = 4,       /* [float32] contributing area                     */
    TF_D8_INMASK = 5,     /* [uint8]   bit k set: the k-th child direction   */
    TF_D8_RAW = 6         /* [int16]   working codes before the uint8 cast   */
};

/* diversion point sets (channels_base.py:1226-1297) */
enum tf_points { TF_PT_SOURCES = 0, TF_PT_SINKS = 1, TF_PT_CANALS_IN = 2,
                 TF_PT_CANALS_OUT = 3 };

/* ---- lifetime ------------------------------------------------------------ */

/* Last error message of the calling thread ("" if none).  The reference prints
 * and raises (channels_base.py:2028, :2078; outlets.py:81); the Python layer
 * turns this text into the same exceptions. */
const char* tf_last_error(void);

/* Library/build identification, e.g. "topoflow_b200 0.1 sm_100a" (the
 * reference's get_attribute('version'), channels_kinematic_wave.py:42). */
const char* tf_version(void);

/* Number of CUDA devices visible, or -1 on error (no compute is started).
 * No counterpart in the reference (CPU only). */
int tf_device_count(void);

/* Create a strip of ny rows x nx columns on CUDA device `device`.
 * Replaces the array allocations spread over channels_component.initialize()
 * (channels_base.py:467-571) and d8_component.initialize_computed_vars()
 * (d8_base.py:505-519).  flags: reserved, pass 0. */
int tf_create(int nx, int ny, int device, unsigned flags, tf_handle* out);

/* Place the strip inside a larger grid of global_ny rows: its first row is
 * global row `row_offset`; has_north / has_south say whether another strip
 * (not the grid border) lies beyond its first / last row.  Call before
 * tf_d8_build.  No counterpart in the reference (single process). */
int tf_set_strip(tf_handle h, int row_offset, int global_ny, int has_north,
                 int has_south);

/* Free every device buffer of the handle (channels_base.py:708-722 finalize). */
int tf_destroy(tf_handle h);

/* Block until all work queued on the handle has finished.  No counterpart in
 * the reference (every NumPy call is synchronous). */
int tf_sync(tf_handle h);

/* ---- D8 grid builder ------------------------------------------------------ */

/* Flow codes, parent IDs, ds, dw and the child mask from a DEM.
 * Replaces d8_component.update_most(): update_flow_grid (d8_global.py:97-144:
 * start_new_d8_codes :179-334, break_flow_grid_ties :352-367, link_flats
 * :371-567), update_parent_ID_grid (:574-644), update_flow_width_grid
 * (:863-935), update_flow_length_grid (:949-1004), update_noflow_IDs
 * (:839-859).
 *   dem        float32 [ny][nx]
 *   dx,dy,dd   float32 [ny] per-row pixel sizes (d8_base.py:559-573)
 *   nodata     cells <= nodata or non-finite get code 0 (d8_base.py:184)
 *   link_flats / lr_periodic / tb_periodic : the cfg switches of the same name
 *   stats_out  optional int64[4]: n_flats, n_pits, n_nodata, link_flats reps */
int tf_d8_build(tf_handle h, const float* dem, const float* dx,
                const float* dy, const float* dd, float nodata,
                int link_flats, int lr_periodic, int tb_periodic,
                int64_t* stats_out);

/* Pixel sizes of row 0 of the GLOBAL grid.  start_new_d8_codes divides every
 * row's elevation drops by the distances of row 0 (d8_global.py:209-211), so a
 * strip that does not hold global row 0 must not use its own first row's on a
 * grid whose dx varies by row (geographic pixels, pixels.py:72-173).  Optional:
 * without this call a strip inside a communicator takes rank 0's values over
 * NCCL in tf_d8_build; a strip with neither uses its own dx[0], dy[0], dd[0]
 * (right for the whole grid and for fixed-length pixels). */
int tf_d8_set_row0_sizes(tf_handle h, float dx0, float dy0, float dd0);

/* Install a ready-made uint8 D8 code grid instead of deriving it from a DEM
 * (what the pre-3.5 tf_d8_base did with <site>_flow.rtg; d8_global.py:162-166).
 * Derives parent IDs, ds, dw and the child mask exactly as tf_d8_build. */
int tf_d8_set_codes(tf_handle h, const uint8_t* codes, const float* dx,
                    const float* dy, const float* dd,
                    int lr_periodic, int tb_periodic);

/* Contributing area.  Replaces d8_component.update_area_grid
 * (d8_global.py:1018-1261).  da_rows: float64 [ny] cell area per row in m^2
 * (pixels.py:174-200); units_km != 0 divides by 1e6 first (A_units 'km^2').
 * n_unresolved_out (optional) counts flowing cells left at 0 (cycles).
 * On a strip with neighbours (tf_set_strip + tf_comm_init) the call is
 * collective.  Flow paths cross strips and the float32 sums must see the true
 * child values in the reference's order, so each strip accumulates what it
 * can, the strips exchange the edge rows of the area grid, every cell whose
 * last pending child lay in a neighbour's row is then computed and its flow
 * path followed on, and so on until no strip has anything new: as many
 * rounds as the most boundary crossings of one flow path, the same bits as
 * one strip, any size of the global grid (da_rows is this strip's [ny]).  A
 * periodic grid (flow wraps around the outer edges) instead gathers the code
 * grid and accumulates the whole grid on every GPU (< 2^31 cells). */
int tf_d8_area(tf_handle h, const double* da_rows, int units_km,
               int64_t* n_unresolved_out);

/* Strips: how many exchange rounds the last tf_d8_area took (0 for a single
 * strip or the replicated route).  No counterpart in the reference. */
long long tf_d8_area_rounds(tf_handle h);

/* Copy one D8 product to a dense host grid of the field's element type: the
 * attributes other components read from the embedded d8 object (d8_grid,
 * parent_ID_grid, ds, dw, A; d8_base.py:119-123, channels_base.py:756-793). */
int tf_d8_get(tf_handle h, int which, void* host_out);

/* ---- channel routing ------------------------------------------------------ */

/* One static parameter, as a float64 grid (grid != NULL) or a scalar
 * broadcast to every cell.  is_grid_in_reference tells whether the reference
 * holds this variable as a grid: only grid-typed width / angle are rewritten
 * by the overbank rule (channels_base.py:1449-1454). */
int tf_chan_set_param(tf_handle h, int which, const double* grid,
                      double scalar, int is_grid_in_reference);

/* The bank angle [rad] as the reference holds it (channels_base.py:2191, "angle
 * *= deg_to_rad"), instead of TF_P_TAN_ANGLE / TF_P_COS_ANGLE grids: the library
 * finds the distinct values on the device and returns them in ascending order
 * (distinct_out: room for 255; *n_out = -1 when there are more, then hand the
 * tan / cos grids over with tf_chan_set_param).  The caller evaluates tan and cos
 * of those few values with ITS libm -- the one the reference's NumPy uses, which
 * is what makes vol and d bit-identical -- and gives them back with
 * tf_chan_set_angle_table (same order).  Saves two transcendental passes and two
 * grid uploads per initialisation. */
int tf_chan_set_angle(tf_handle h, const double* grid, double scalar, int is_grid_in_reference,
                      double* distinct_out, int* n_out);
int tf_chan_set_angle_table(tf_handle h, const double* tan_values, const double* cos_values, int n);

/* Fix the run configuration and build the device layout.  Call after
 * tf_d8_build/tf_d8_set_codes and after all six tf_chan_set_param calls.
 * Replaces set_computed_input_vars (channels_base.py:726-752) and the rest of
 * initialize_computed_vars (:797-984).
 *   wave        enum tf_wave  (cfg_extension, :732-736)
 *   law         enum tf_law   (MANNING / LAW_OF_WALL cfg switches)
 *   dt          channel time step [s]
 *   da_rows     float64 [ny] cell area per row [m^2]
 *   outlet_rc   int32 [2*n_outlets] (row, col) pairs; pair 0 is outlet_ID
 *   allow_packed  != 0 lets the library store parameters that are exactly
 *                 representable in float32 as float32 and bank angles as a
 *                 table index; results are unchanged, traffic is lower. */
int tf_chan_init(tf_handle h, int wave, int law, double dt,
                 const double* da_rows, const int32_t* outlet_rc,
                 int n_outlets, int allow_packed);

/* update_R + update_R_integral on the device (channels_base.py:1103-1160).
 * tf_chan_set_forcing hands over one of the six fluxes the framework sets
 * into the component every step, as a dense [ny, nx] float64 grid or (grid ==
 * NULL) a scalar; tf_chan_assemble_R then forms
 *     R = (P_rain + SM + GW + MR) - (ET + IN)
 * cell by cell in that order of operations (a scalar flux broadcasts as in
 * NumPy), leaves it where the step kernels read it, and returns in sum_out
 * (optional) the vol_R increment sum((R*da)*dt).  With all six scalar, R
 * stays a scalar kernel argument and no grid is touched. */
int tf_chan_set_forcing(tf_handle h, int which, const double* grid, double scalar);
int tf_chan_assemble_R(tf_handle h, double* sum_out);

/* Density of water [kg m-3] used for tau and u_star (the met component's
 * 'water-liquid__mass-per-volume_density', channels_base.py:165, :1577). */
int tf_chan_set_rho(tf_handle h, double rho);

/* Overwrite one materialised state grid (TF_F_VOL, TF_F_D, TF_F_U, TF_F_Q, TF_F_QNEXT)
 * with host data: the initial state of initialize_computed_vars
 * (channels_base.py:900-960), or a restart. */
int tf_chan_set_state(tf_handle h, int which, const double* grid);

/* Excess rain rate R = (P_rain + SM + GW + MR) - (ET + IN) [m/s]
 * (channels_base.py:1103-1147), as a grid or a scalar.  sum_out (optional)
 * receives sum over the strip of (R*da)*dt, the per-step increment of vol_R
 * (update_R_integral, :1151-1160). */
int tf_chan_set_R(tf_handle h, const double* grid, double scalar,
                  double* sum_out);

/* Diversion point data (channels_base.py:1226-1297).  rows/cols int32[n];
 * values float64[n]: sources_Q, sinks_Q, canals_Q_fraction or canals_out_Q. */
int tf_chan_set_points(tf_handle h, int kind, const int32_t* rows,
                       const int32_t* cols, const double* values, int n);

/* Advance n_steps time steps: update_discharge .. check_flow_velocity of
 * channels_component.update() (channels_base.py:602-672) per step.
 *   taps_out      float64 [n_steps][n_outlets][4] = Q, u, d, f at the outlets
 *                 after each step (update_outlet_values, :1812-1832)
 *   canals_in_out float64 [n_steps][n_canals] canals_in_Q (:1283), or NULL
 *   bad_out       int64[2]: first cell (row-major index) with a negative or
 *                 non-finite depth, then velocity, -1 if none
 *                 (check_flow_depth :1982-2030, check_flow_velocity :2034-2080)
 * Host buffers; the copies are part of the call. */
int tf_chan_step(tf_handle h, int n_steps, double* taps_out,
                 double* canals_in_out, int64_t* bad_out);

/* Same work, nothing copied back: queue n_steps on the handle's stream and
 * return.  Used to time the kernels with inputs resident in HBM.  No counterpart
 * in the reference. */
int tf_chan_step_async(tf_handle h, int n_steps);

/* Copy one grid (enum tf_field) to a dense float64 host grid.  Derived fields
 * are recomputed from the state the way update_trapezoid_Rh (:1590-1643),
 * update_shear_stress/speed (:1567-1586), update_friction_factor (:1654-1721)
 * and update_froude_number (:1795-1808) do. */
int tf_chan_get(tf_handle h, int which, double* host_out);

/* Interior min/max of Q, u, d: out[6] = Qmin,Qmax,umin,umax,dmin,dmax
 * (update_mins_and_maxes, channels_base.py:1890-1916). */
int tf_chan_minmax(tf_handle h, double* out6);

/* ---- output frames ---------------------------------------------------------
 * Replaces the grid branch of write_output_files (channels_base.py:2465-2535)
 * -> model_output.add_grid -> rts_file.add_grid (utils/rts_files.py:313-396):
 * "grid = np.float32(grid); if SWAP_ENDIAN: grid.byteswap(True); tofile".
 * The field (enum tf_field) is rounded to float32 (round to nearest even, as
 * np.float32 does) and, when swap_endian != 0, byte-swapped ON THE DEVICE into
 * a dense [ny, nx] frame, which is copied to host_out on a separate copy
 * stream.  async != 0 returns at once: later tf_chan_step* calls overlap the
 * copy, and host_out (best allocated with tf_host_alloc) is valid after
 * tf_chan_frame_wait().  One frame may be in flight per handle; a second call
 * waits on the device for the first copy to leave the frame buffer. */
int tf_chan_frame(tf_handle h, int which, int swap_endian, float* host_out, int async);
int tf_chan_frame_wait(tf_handle h);

/* Page-locked host memory for frames and forcing grids (cudaHostAlloc).  No
 * counterpart in the reference; it replaces the NumPy buffers of
 * rts_file.add_grid (utils/rts_files.py:313-396) on the caller's side. */
int tf_host_alloc(size_t bytes, void** out);
int tf_host_free(void* p);

/* (No counterpart in the reference beyond the cProfile notes in its comments,
 * channels_base.py:1398-1400.)
 * Device-side timing of everything queued by tf_chan_step* since the last
 * call: out[0] = elapsed ms (CUDA events on the handle's stream), out[1] =
 * steps, out[2] = kernel launches, out[3] = bytes per cell-update actually
 * addressed by the step kernels for the current layout. */
int tf_chan_timing(tf_handle h, double* out4);

/* sum over an ny x nx device grid of (R*da)*dt -- the vol_R increment that
 * tf_chan_set_R returns -- for callers that are not allowed to allocate
 * device memory: code inside a stream capture, plug-ins whose host owns all
 * GPU memory through its own pool, loops that cannot take the device-wide
 * stall a cudaMalloc/cudaFree can cause.  This entry point has no counterpart
 * in the reference; it exists only for such callers.
 *   R_dev        device pointer, row y starts at R_dev + y*pitch (elements)
 *   da_rows_dev  device float64 [ny] cell area per row, or NULL to use `da`
 *   sum_out_dev  device pointer to one float64, written when the stream runs
 *   stream       a cudaStream_t, or NULL for the default stream
 * It calls no cudaMalloc / cudaFree and no allocating library routine.  Its
 * per-row partial sums live in a fixed array inside the library image
 * (65536 rows x 8 B = 512 KiB per GPU, held until the process exits); a grid
 * with more rows is refused with an error -- use tf_chan_set_R for those.
 * The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that first call before any stream capture
 * starts.  All calls on a device share the one array: two calls whose kernels
 * can overlap on the GPU (different streams) race on it, so keep them on one
 * stream or order them with events.  Nothing has to be zeroed between calls:
 * every slot that is read is overwritten first. */
int tf_sum_R_noalloc(const double* R_dev, int nx, int ny, long long pitch,
                     const double* da_rows_dev, double da, double dt,
                     double* sum_out_dev, void* stream);

/* ---- device-resident providers of the excess-rain fluxes --------------------
 * The six fluxes of update_R come, in the reference, from the meteorology,
 * snow, evaporation and infiltration components; EMELI copies their outputs
 * into the channels component through host arrays at every channel step
 * (framework/emeli.py:1841-1933, time_interpolation.py:574-685).  The entry
 * points below run the same expressions on grids that stay on the device.
 * Every operand is given as (grid id, scalar): id >= 0 names a device grid made
 * by tf_grid_new, id < 0 means "use the scalar" -- the reference's Grid /
 * Scalar input types.  Grids are float64 [ny][nx] like everything else. */

/* A new zero-filled device grid; its id in *gid_out.  (The arrays the provider
 * components allocate in their initialize_computed_vars, e.g.
 * snow_base.py:300-345, infil_green_ampt.py:330-360.)  Freed with the handle. */
int tf_grid_new(tf_handle h, int* gid_out);
/* Fill a device grid from a host grid (host != NULL) or with a scalar. */
int tf_grid_set(tf_handle h, int gid, const double* host, double scalar);
/* Copy a device grid to the host (BMI get_values of a provider's variable). */
int tf_grid_get(tf_handle h, int gid, double* host);
/* dst = src on the device. */
int tf_grid_copy(tf_handle h, int src_gid, int dst_gid);
/* out = a * t + b: the value EMELI hands a user at time t
 * (time_interpolation.py:677). */
int tf_grid_axpb(tf_handle h, int a_gid, double a, int b_gid, double b, double t, int out_gid);
/* After a provider's update: the (a, b) of EMELI's linear interpolation
 * between the previous value v1 (time t1) and the new one v2 (time t2),
 * time_interpolation.py:91-133 -- a = (v2 - v1) / (t2 - t1), b = v2 - a t2 when
 * EVERY cell changed (min |v2 - v1| != 0, over all strips) and t2 != t1, else
 * a = 0, b = v2.  *linear_out (optional) tells which. */
int tf_interp_linear(tf_handle h, int v1_gid, int v2_gid, double t1, double t2,
                     int a_gid, int b_gid, int* linear_out);

/* P_rain = P * (T_air > 0), P_snow = P * (T_air <= 0)  (met_base.py:1013-1038,
 * :1060-1085).  Either output id may be < 0. */
int tf_prov_met_rain(tf_handle h, int P_gid, double P, int T_air_gid, double T_air,
                     int P_rain_gid, int P_snow_gid);
/* One update of the degree-day snow component: meltrate
 * (snow_degree_day.py:272-312), its cap by the snow present and the
 * water-equivalent / depth bookkeeping (snow_base.py:365-440).  h_snow, h_swe
 * are state grids updated in place; SM is written; *vol_SM_out (optional) gets
 * sum((SM*da)*dt) (update_SM_integral).  Call after tf_chan_init (cell areas). */
int tf_prov_snow_degree_day(tf_handle h, int T_air_gid, double T_air, int P_snow_gid, double P_snow,
                            int c0_gid, double c0, int T0_gid, double T0, double rho_H2O,
                            int rho_snow_gid, double rho_snow, double dt,
                            int h_snow_gid, int h_swe_gid, int SM_gid, double* vol_SM_out);
/* Priestley-Taylor evaporation rate (evap_priestley_taylor.py:290-353):
 * ET = max(alpha (0.406 + 0.011 T_air) (Qn_SW + Qn_LW - Qc) / 2.5e9, 0),
 * Qc = K_soil (T_soil_x - T_surf) / soil_x.  *vol_ET_out: sum((ET*da)*dt). */
int tf_prov_evap_priestley_taylor(tf_handle h, int Qn_SW_gid, double Qn_SW, int Qn_LW_gid, double Qn_LW,
                                  int T_air_gid, double T_air, int T_surf_gid, double T_surf,
                                  int alpha_gid, double alpha, int K_soil_gid, double K_soil,
                                  int soil_x_gid, double soil_x, int T_soil_x_gid, double T_soil_x,
                                  double dt, int ET_gid, double* vol_ET_out);
/* One update of the one-layer Green-Ampt infiltration component
 * (infil_base.py:216-264): P_total = P_rain + SM (:500-513), IN = min(G dK dq /
 * I + Ks, P_total) (infil_green_ampt.py:516-572), IN = P_total where P_total <
 * Ks (infil_base.py:784-823), IN = 0 where h_table >= elev when both grids are
 * given (:592-647), I += IN dt (:700-728).  I is state, IN is written. */
int tf_prov_infil_green_ampt(tf_handle h, int P_rain_gid, double P_rain, int SM_gid, double SM,
                             int Ks_gid, double Ks, int Ki_gid, double Ki, int qs_gid, double qs,
                             int qi_gid, double qi, int G_gid, double G, int h_table_gid, int elev_gid,
                             double dt, int I_gid, int IN_gid, double* vol_IN_out);

/* Flux `which` (enum tf_forcing) of update_R as value + rate * t, both device
 * grids or scalars: the (b, a) pair of tf_interp_linear, or (value, 0).
 * tf_chan_assemble_R_dev then forms R0 = (P_rain + SM + GW + MR) - (ET + IN)
 * of the values (channels_base.py:1103-1147, in that order) and R1 = the same
 * combination of the rates; the step kernels use R = R1 * t + R0 with t the
 * channel time of each step.  Nothing crosses PCIe.  sum0_out / sum1_out
 * (optional): sum((R0*da)*dt) and sum((R1*da)*dt), from which the caller keeps
 * update_R_integral (:1151-1160) as sum1 * t + sum0 per step. */
int tf_chan_set_forcing_dev(tf_handle h, int which, int value_gid, double value,
                            int rate_gid, double rate);
int tf_chan_assemble_R_dev(tf_handle h, double* sum0_out, double* sum1_out);

/* ---- neighbours of the path on the same D8 structures ----------------------- */

/* Depression filling of a float32 DEM: the result of utils/fill_pits.py
 * fill_pits() (:331-582, Wang & Liu's priority flood from the start cells of
 * get_start_pixels, :178-330: edge cells and cells next to nodata).  Every cell
 * ends at max(z, lowest pass on any path to a start cell); non-finite cells
 * become nodata, nodata cells are walls.  The serial heap is replaced by a
 * relaxation from above that reaches the same, unique surface (bit-identical:
 * elevations are only ever copied).  One strip only (tf_set_strip not used).
 *   n_raised_out  optional: cells whose elevation was raised ("Raised pixels")
 *   n_sweeps_out  optional: passes over the grid it took */
int tf_fill_pits(tf_handle h, const float* dem, float nodata, float* filled_out,
                 int64_t* n_raised_out, int64_t* n_sweeps_out);

/* The lateral-flow part of satzone_base.update_water_table
 * (components/satzone_base.py:930-1008): dzw = dt (Rg - Q_gw / da), then for k =
 * 1..8 dzw[p_k] += Q_gw[w_k] dt / da -- the scatter of update_flow_volume on
 * the ground-water discharge, as a gather over the child mask in the same
 * order.  Q_gw, dzw_out: float64 [ny][nx]; Rg: grid or (NULL) scalar; da_rows:
 * float64 [ny] cell area per row (a uniform da takes the reference's
 * "factor = dt / da" branch, :976-992, a varying one the other, :993-1008).
 * Needs the D8 grid (tf_d8_build). */
int tf_satzone_dzw(tf_handle h, const double* Q_gw, const double* Rg_grid, double Rg_scalar,
                   const double* da_rows, double dt, double* dzw_out);

/* ---- row-strip decomposition over several GPUs ---------------------------- */

/* 128-byte NCCL unique id for tf_comm_init (rank 0 creates, all ranks use).
 * The reference is a single process: nothing in it corresponds to tf_comm_*. */
int tf_comm_unique_id(uint8_t* id128);

/* Join a communicator of `nranks` strips stacked north to south; rank r's
 * north neighbour is r-1.  After this, tf_chan_step* exchanges one halo row
 * of Q (plus d and u for the diffusive / dynamic wave) with each neighbour
 * per step over NCCL send/recv.  The reference has no counterpart: it is
 * single-process (docs/Peckham_2014_EMELI_FINAL.pdf). */
int tf_comm_init(tf_handle h, int rank, int nranks, const uint8_t* id128);

/* Which fields (enum tf_field, -1 terminated, fields_out3[3]) a strip exchanges
 * with its neighbours in one time step of wave `wave`: point 0 = between the
 * volume/depth phase and the velocity phase, point 1 = at the end of the step.
 * The launch sequence of tf_chan_step reads the same table.  No counterpart in
 * the reference (single process). */
int tf_halo_plan(int wave, int point, int* fields_out3);

/* Set the halo DEM rows by hand (single process tests): north/south float32
 * [nx] or NULL for "no neighbour" (global border). */
int tf_d8_set_halo_dem(tf_handle h, const float* north_row,
                       const float* south_row);

#ifdef __cplusplus
}
#endif
#endif /* TOPOFLOW_B200_H */
