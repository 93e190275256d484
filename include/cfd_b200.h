/*
 * cfd_b200.h -- C-ABI of libcfd_b200.so, the B200-native (sm_100a) backend for the
 * shaia/CFD projection-method time step.
 *
 * Part 1 mirrors, field for field, the reference's own ABI types that cross the
 *        boundary (so this header is usable without the reference tree).
 * Part 2 declares the drop-in entry points: the very symbols that the reference's
 *        solver registry and Poisson dispatcher bind at link time
 *        (lib/src/api/solver_registry.c:83-95, lib/src/solvers/linear/linear_solver.c:165-229).
 * Part 3 is the backend-private extension surface (cfd_b200_*): Poisson-method
 *        selection (the reference has no public knob), device-resident sessions for
 *        multi-step / multi-GPU runs, and measurement hooks.
 *
 * All signatures are plain C: pointers, sizes, doubles and ints. No C++/torch types.
 * When the real reference headers were already included, Part 1 is skipped.
 */
#ifndef CFD_B200_H
#define CFD_B200_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CFD_B200_API __attribute__((visibility("default")))
#else
#define CFD_B200_API
#endif

/* ======================================================================
 * Part 1 -- ABI mirror of the reference types
 * ====================================================================== */
#ifndef CFD_NAVIER_STOKES_SOLVER_H /* reference: lib/include/cfd/solvers/navier_stokes_solver.h */

/* lib/include/cfd/core/cfd_status.h:13-24 */
typedef enum {
    CFD_SUCCESS = 0,
    CFD_ERROR = -1,
    CFD_ERROR_NOMEM = -2,
    CFD_ERROR_INVALID = -3,
    CFD_ERROR_IO = -4,
    CFD_ERROR_UNSUPPORTED = -5,
    CFD_ERROR_DIVERGED = -6,
    CFD_ERROR_MAX_ITER = -7,
    CFD_ERROR_LIMIT_EXCEEDED = -8,
    CFD_ERROR_NOT_FOUND = -9
} cfd_status_t;

/* lib/include/cfd/core/grid.h:18-40 */
typedef struct {
    double* x;
    double* y;
    double* dx;
    double* dy;
    size_t nx;
    size_t ny;
    double xmin, xmax, ymin, ymax;
    double* z;
    double* dz;
    size_t nz;
    double zmin, zmax;
    size_t stride_z;
    double inv_dz2;
    size_t k_start;
    size_t k_end;
} grid;

/* lib/include/cfd/boundary/boundary_conditions.h:19-27 */
typedef enum {
    BC_TYPE_PERIODIC,
    BC_TYPE_NEUMANN,
    BC_TYPE_DIRICHLET,
    BC_TYPE_NOSLIP,
    BC_TYPE_INLET,
    BC_TYPE_OUTLET,
    BC_TYPE_SYMMETRY
} bc_type_t;

/* lib/include/cfd/boundary/boundary_conditions.h:50-57 */
typedef struct {
    double left, right, top, bottom, front, back;
} bc_dirichlet_values_t;

/* boundary_conditions.h:64-68, 85-89, 96-103: inlet profile / specification / edge */
typedef enum { BC_INLET_PROFILE_UNIFORM, BC_INLET_PROFILE_PARABOLIC, BC_INLET_PROFILE_CUSTOM } bc_inlet_profile_t;
typedef enum { BC_INLET_SPEC_VELOCITY, BC_INLET_SPEC_MAGNITUDE_DIR, BC_INLET_SPEC_MASS_FLOW } bc_inlet_spec_type_t;
typedef enum {
    BC_EDGE_LEFT = 0x01,
    BC_EDGE_RIGHT = 0x02,
    BC_EDGE_BOTTOM = 0x04,
    BC_EDGE_TOP = 0x08,
    BC_EDGE_FRONT = 0x10,
    BC_EDGE_BACK = 0x20
} bc_edge_t;

/* boundary_conditions.h:117-198: time variation of an inlet (carried in bc_inlet_config_t; the
 * steady entry point bc_apply_inlet_gpu does not evaluate it, as in the reference) */
typedef enum {
    BC_TIME_PROFILE_CONSTANT,
    BC_TIME_PROFILE_SINUSOIDAL,
    BC_TIME_PROFILE_RAMP,
    BC_TIME_PROFILE_STEP,
    BC_TIME_PROFILE_CUSTOM
} bc_time_profile_t;
typedef struct { double frequency, amplitude, phase, offset; } bc_time_sinusoidal_t;
typedef struct { double t_start, t_end, value_start, value_end; } bc_time_ramp_t;
typedef struct { double t_step, value_before, value_after; } bc_time_step_t;
typedef double (*bc_time_custom_fn)(double time, double dt, void* user_data);
typedef struct {
    bc_time_profile_t profile;
    union {
        bc_time_sinusoidal_t sinusoidal;
        bc_time_ramp_t ramp;
        bc_time_step_t step;
    } params;
    bc_time_custom_fn custom_fn;
    void* custom_user_data;
} bc_time_config_t;
typedef void (*bc_inlet_profile_fn)(double position, double* u_out, double* v_out, void* user_data);
typedef void (*bc_inlet_profile_time_fn)(double position, double time, double dt, double* u_out, double* v_out,
                                         void* user_data);

/* boundary_conditions.h:233-267 */
typedef struct {
    bc_edge_t edge;
    bc_inlet_profile_t profile;
    bc_inlet_spec_type_t spec_type;
    union {
        struct { double u, v; } velocity;
        struct { double magnitude, direction; } magnitude_dir;
        struct { double mass_flow_rate, density, inlet_length; } mass_flow;
    } spec;
    bc_inlet_profile_fn custom_profile;
    void* custom_profile_user_data;
    bc_time_config_t time_config;
    bc_inlet_profile_time_fn custom_profile_time;
    void* custom_profile_time_user_data;
} bc_inlet_config_t;

/* lib/include/cfd/solvers/navier_stokes_solver.h:54-64 */
typedef struct {
    double* u;
    double* v;
    double* w;
    double* p;
    double* rho;
    double* T;
    size_t nx, ny, nz;
} flow_field;

/* navier_stokes_solver.h:77-92 */
typedef void (*ns_source_func_t)(double x, double y, double z, double t, void* context,
                                 double* source_u, double* source_v, double* source_w);
typedef double (*ns_heat_source_func_t)(double x, double y, double z, double t, void* context);

/* navier_stokes_solver.h:108-116 */
typedef struct {
    bc_type_t left, right, bottom, top, front, back;
    bc_dirichlet_values_t dirichlet_values;
} ns_thermal_bc_config_t;

/* navier_stokes_solver.h:121-158 */
typedef struct {
    double dt, cfl, gamma, mu, k;
    int max_iter;
    double tolerance;
    double source_amplitude_u, source_amplitude_v, source_decay_rate, pressure_coupling;
    ns_source_func_t source_func;
    void* source_context;
    double alpha, beta, T_ref;
    double gravity[3];
    ns_heat_source_func_t heat_source_func;
    void* heat_source_context;
    ns_thermal_bc_config_t thermal_bc;
} ns_solver_params_t;

/* navier_stokes_solver.h:172-207 */
typedef enum {
    NS_SOLVER_BACKEND_SCALAR = 0,
    NS_SOLVER_BACKEND_SIMD = 1,
    NS_SOLVER_BACKEND_OMP = 2,
    NS_SOLVER_BACKEND_CUDA = 3
} ns_solver_backend_t;

typedef enum {
    NS_SOLVER_CAP_NONE = 0,
    NS_SOLVER_CAP_INCOMPRESSIBLE = (1 << 0),
    NS_SOLVER_CAP_COMPRESSIBLE = (1 << 1),
    NS_SOLVER_CAP_STEADY_STATE = (1 << 2),
    NS_SOLVER_CAP_TRANSIENT = (1 << 3),
    NS_SOLVER_CAP_SIMD = (1 << 4),
    NS_SOLVER_CAP_PARALLEL = (1 << 5),
    NS_SOLVER_CAP_GPU = (1 << 6)
} ns_solver_capabilities_t;

typedef struct {
    int iterations;
    double residual;
    double max_velocity;
    double max_pressure;
    double max_temperature;
    double cfl_number;
    double elapsed_time_ms;
    cfd_status_t status;
} ns_solver_stats_t;

/* navier_stokes_solver.h:213-277 -- the solver vtable a registry factory returns */
typedef struct NSSolver ns_solver_t;
typedef void* ns_solver_context_t;
typedef cfd_status_t (*ns_solver_init_func)(ns_solver_t*, const grid*, const ns_solver_params_t*);
typedef void (*ns_solver_destroy_func)(ns_solver_t*);
typedef cfd_status_t (*ns_solver_step_func)(ns_solver_t*, flow_field*, const grid*,
                                            const ns_solver_params_t*, ns_solver_stats_t*);
typedef cfd_status_t (*ns_solver_solve_func)(ns_solver_t*, flow_field*, const grid*,
                                             const ns_solver_params_t*, ns_solver_stats_t*);
typedef void (*ns_solver_boundary_func)(ns_solver_t*, flow_field*, const grid*);
typedef double (*ns_solver_compute_dt_func)(ns_solver_t*, const flow_field*, const grid*,
                                            const ns_solver_params_t*);
typedef const char* (*ns_solver_get_name_func)(const ns_solver_t*);
typedef const char* (*ns_solver_get_description_func)(const ns_solver_t*);
typedef ns_solver_capabilities_t (*ns_solver_get_capabilities_func)(const ns_solver_t*);

struct NSSolver {
    const char* name;
    const char* description;
    const char* version;
    ns_solver_capabilities_t capabilities;
    ns_solver_backend_t backend;
    ns_solver_context_t context;
    ns_solver_init_func init;
    ns_solver_destroy_func destroy;
    ns_solver_step_func step;
    ns_solver_solve_func solve;
    ns_solver_boundary_func apply_boundary;
    ns_solver_compute_dt_func compute_dt;
    ns_solver_get_name_func get_name;
    ns_solver_get_description_func get_description;
    ns_solver_get_capabilities_func get_capabilities;
};

typedef ns_solver_t* (*ns_solver_factory_func)(void);

#endif /* CFD_NAVIER_STOKES_SOLVER_H */

#ifndef CFD_POISSON_SOLVER_H /* reference: lib/include/cfd/solvers/poisson_solver.h */

/* poisson_solver.h:53-108 */
typedef enum {
    POISSON_METHOD_JACOBI,
    POISSON_METHOD_GAUSS_SEIDEL,
    POISSON_METHOD_SOR,
    POISSON_METHOD_REDBLACK_SOR,
    POISSON_METHOD_CG,
    POISSON_METHOD_BICGSTAB,
    POISSON_METHOD_MULTIGRID
} poisson_solver_method_t;

typedef enum {
    POISSON_BACKEND_AUTO,
    POISSON_BACKEND_SCALAR,
    POISSON_BACKEND_OMP,
    POISSON_BACKEND_SIMD,
    POISSON_BACKEND_GPU
} poisson_solver_backend_t;

typedef enum {
    POISSON_CONVERGED = 0,
    POISSON_MAX_ITER = 1,
    POISSON_DIVERGED = 2,
    POISSON_STAGNATED = 3,
    POISSON_ERROR = -1
} poisson_solver_status_t;

typedef enum { POISSON_PRECOND_NONE = 0, POISSON_PRECOND_JACOBI = 1 } poisson_precond_type_t;
/* No counterpart in the reference (its roadmap item ROADMAP.md:190-194): POISSON_METHOD_CG preconditioned by one
 * multigrid V(1,1) cycle built for the Krylov operator itself (zero boundary cells, csrc/poisson_mgpc.cu).  Same
 * system and stopping rule as the reference's CG, 7-8 iterations instead of hundreds; the iterates differ, the
 * answers agree to the tolerance asked for.  One GPU. */
#define CFD_B200_PRECOND_MULTIGRID ((poisson_precond_type_t)100)

typedef struct {
    double tolerance;
    double absolute_tolerance;
    int max_iterations;
    double omega;
    int check_interval;
    bool verbose;
    poisson_precond_type_t preconditioner;
} poisson_solver_params_t;

typedef struct {
    poisson_solver_status_t status;
    int iterations;
    double initial_residual;
    double final_residual;
    double elapsed_time_ms;
} poisson_solver_stats_t;

/* poisson_solver.h:163-233 -- the Poisson vtable a factory returns */
typedef struct poisson_solver poisson_solver_t;
typedef void* poisson_solver_context_t;
typedef cfd_status_t (*poisson_solver_init_func)(poisson_solver_t*, size_t nx, size_t ny, size_t nz,
                                                 double dx, double dy, double dz,
                                                 const poisson_solver_params_t*);
typedef void (*poisson_solver_destroy_func)(poisson_solver_t*);
typedef cfd_status_t (*poisson_solver_solve_func)(poisson_solver_t*, double* x, double* x_temp,
                                                  const double* rhs, poisson_solver_stats_t*);
typedef cfd_status_t (*poisson_solver_iterate_func)(poisson_solver_t*, double* x, double* x_temp,
                                                    const double* rhs, double* residual);
typedef void (*poisson_solver_apply_bc_func)(poisson_solver_t*, double* x);

struct poisson_solver {
    const char* name;
    const char* description;
    poisson_solver_method_t method;
    poisson_solver_backend_t backend;
    size_t nx, ny, nz;
    double dx, dy, dz;
    poisson_solver_params_t params;
    poisson_solver_context_t context;
    poisson_solver_init_func init;
    poisson_solver_destroy_func destroy;
    poisson_solver_solve_func solve;
    poisson_solver_iterate_func iterate;
    poisson_solver_apply_bc_func apply_bc;
};

#endif /* CFD_POISSON_SOLVER_H */

#ifndef CFD_GPU_DEVICE_H /* reference: lib/include/cfd/core/gpu_device.h */

/* gpu_device.h:32-53 */
typedef struct {
    int enable_gpu;
    size_t min_grid_size;
    int min_steps;
    int block_size_x;
    int block_size_y;
    int poisson_max_iter;
    double poisson_tolerance;
    int persistent_memory;
    int async_transfers;
    int sync_after_kernel;
    int verbose;
} gpu_config_t;

/* gpu_device.h:58-69 */
typedef struct {
    int device_id;
    char name[256];
    size_t total_memory;
    size_t free_memory;
    int compute_capability_major;
    int compute_capability_minor;
    int multiprocessor_count;
    int max_threads_per_block;
    int warp_size;
    int is_available;
} gpu_device_info_t;

/* gpu_device.h:74-82 */
typedef struct {
    double kernel_time_ms;
    double transfer_time_ms;
    double poisson_time_ms;
    int poisson_iterations;
    double poisson_residual;
    size_t memory_allocated;
    int kernels_launched;
} gpu_solver_stats_t;

typedef struct gpu_solver_context_t gpu_solver_context_t; /* gpu_device.h:88, opaque */

#endif /* CFD_GPU_DEVICE_H */

/* ======================================================================
 * Part 2 -- drop-in entry points (reference names, reference signatures)
 * ====================================================================== */

/* The ten symbols libcfd_api.a needs from the GPU library. */
CFD_B200_API int gpu_is_available(void);                        /* gpu_device.h:103 */
CFD_B200_API gpu_config_t gpu_config_default(void);             /* gpu_device.h:97  */
CFD_B200_API cfd_status_t solve_projection_method_gpu(          /* gpu_device.h:221 */
    flow_field* field, const grid* grid, const ns_solver_params_t* params, const gpu_config_t* config);
CFD_B200_API cfd_status_t solve_rk2_method_gpu(                 /* gpu_device.h:234 */
    flow_field* field, const grid* grid, const ns_solver_params_t* params, const gpu_config_t* config);
CFD_B200_API cfd_status_t solve_rk4_method_gpu(                 /* gpu_device.h:247 */
    flow_field* field, const grid* grid, const ns_solver_params_t* params, const gpu_config_t* config);
CFD_B200_API cfd_status_t solve_explicit_euler_method_gpu(      /* solver_registry.c:86-88 */
    flow_field* field, const grid* grid, const ns_solver_params_t* params, const gpu_config_t* config);
CFD_B200_API poisson_solver_t* create_jacobi_gpu_solver(void);   /* linear_solver.c:165-168 */
CFD_B200_API poisson_solver_t* create_redblack_gpu_solver(void); /* linear_solver.c:194-197 */
CFD_B200_API poisson_solver_t* create_cg_gpu_solver(void);       /* linear_solver.c:212-215 */
CFD_B200_API poisson_solver_t* create_bicgstab_gpu_solver(void); /* linear_solver.c:226-229 */

/* GPU boundary-condition API on DEVICE pointers (cfd/boundary/boundary_conditions_gpu.cuh).
 * `stream` is a cudaStream_t (void* when no CUDA header was included before this one).  Results
 * are those of the reference's CPU passes (x, then y, then z faces), race-free; types other than
 * PERIODIC act as NEUMANN in the generic entry points, as in the reference. */
#if defined(__DRIVER_TYPES_H__) || defined(__CUDA_RUNTIME_H__)
typedef cudaStream_t cfd_b200_stream_t;
#else
typedef void* cfd_b200_stream_t;
#endif
CFD_B200_API void bc_apply_neumann_gpu(double* d_field, size_t nx, size_t ny, cfd_b200_stream_t stream); /* :32 */
CFD_B200_API void bc_apply_scalar_gpu(double* d_field, size_t nx, size_t ny, bc_type_t type, cfd_b200_stream_t stream); /* :43 */
CFD_B200_API void bc_apply_velocity_gpu(double* d_u, double* d_v, size_t nx, size_t ny, bc_type_t type,
                                        cfd_b200_stream_t stream);                                                 /* :55 */
CFD_B200_API void bc_apply_dirichlet_scalar_gpu(double* d_field, size_t nx, size_t ny,
                                                const bc_dirichlet_values_t* values, cfd_b200_stream_t stream);    /* :73 */
CFD_B200_API void bc_apply_dirichlet_velocity_gpu(double* d_u, double* d_v, size_t nx, size_t ny,
                                                  const bc_dirichlet_values_t* u_values,
                                                  const bc_dirichlet_values_t* v_values, cfd_b200_stream_t stream); /* :88 */
CFD_B200_API cfd_status_t bc_apply_inlet_gpu(double* d_u, double* d_v, size_t nx, size_t ny,
                                             const bc_inlet_config_t* config, cfd_b200_stream_t stream);           /* :107 */
CFD_B200_API void bc_apply_scalar_3d_gpu(double* d_field, size_t nx, size_t ny, size_t nz, bc_type_t type,
                                         cfd_b200_stream_t stream);                                                /* :129 */
CFD_B200_API void bc_apply_velocity_3d_gpu(double* d_u, double* d_v, double* d_w, size_t nx, size_t ny, size_t nz,
                                           bc_type_t type, cfd_b200_stream_t stream);                              /* :149 */

/* Remaining public GPU device surface used by the reference's tests/examples. */
CFD_B200_API int gpu_get_device_info(gpu_device_info_t* info, int max_devices); /* gpu_device.h:111 */
CFD_B200_API cfd_status_t gpu_select_device(int device_id);                     /* gpu_device.h:118 */
CFD_B200_API int gpu_should_use(const gpu_config_t* config, size_t nx, size_t ny, size_t nz,
                                int num_steps);                                 /* gpu_device.h:129 */
CFD_B200_API cfd_status_t solve_navier_stokes_gpu(                              /* gpu_device.h:209 */
    flow_field* field, const grid* grid, const ns_solver_params_t* params, const gpu_config_t* config);

/* Legacy persistent-context API (gpu_device.h:145-194). step() advances one projection step. */
CFD_B200_API gpu_solver_context_t* gpu_solver_create(size_t nx, size_t ny, size_t nz,
                                                     const gpu_config_t* config);
CFD_B200_API void gpu_solver_destroy(gpu_solver_context_t* ctx);
CFD_B200_API cfd_status_t gpu_solver_upload(gpu_solver_context_t* ctx, const flow_field* field);
CFD_B200_API cfd_status_t gpu_solver_download(gpu_solver_context_t* ctx, flow_field* field);
CFD_B200_API cfd_status_t gpu_solver_step(gpu_solver_context_t* ctx, const grid* grid,
                                          const ns_solver_params_t* params, gpu_solver_stats_t* stats);
CFD_B200_API gpu_solver_stats_t gpu_solver_get_stats(const gpu_solver_context_t* ctx);
CFD_B200_API void gpu_solver_reset_stats(gpu_solver_context_t* ctx);

/* Run-time registration path (navier_stokes_solver.h:333-337): factories returning a heap
 * ns_solver_t with real init/destroy (persistent device context) and fully filled stats. */
CFD_B200_API ns_solver_t* cfd_b200_create_projection_gpu_solver(void);
CFD_B200_API ns_solver_t* cfd_b200_create_explicit_euler_gpu_solver(void); /* solver_registry.c:1247-1291 */
CFD_B200_API ns_solver_t* cfd_b200_create_rk2_gpu_solver(void);
CFD_B200_API ns_solver_t* cfd_b200_create_rk4_gpu_solver(void);
/* Geometric multigrid behind the poisson_solver_t vtable (the reference reserves POISSON_METHOD_MULTIGRID,
 * poisson_solver.h:60, without an implementation; ROADMAP.md:116,194).  iterations = V-cycles, stopping rule
 * of linear_solver.c:397-485 on the L-infinity residual; default cap 100 cycles.  One GPU only. */
CFD_B200_API poisson_solver_t* cfd_b200_create_multigrid_gpu_solver(void);

/* ======================================================================
 * Part 3 -- backend-private extensions
 * ====================================================================== */

/* What to do when the Poisson solve hits its iteration cap. */
typedef enum {
    CFD_B200_NONCONV_ERROR = 0,   /* reference CPU behaviour: CFD_ERROR_MAX_ITER (solver_projection.c:220-224) */
    CFD_B200_NONCONV_CONTINUE = 1 /* reference GPU behaviour / fixed-iteration timing mode: proceed */
} cfd_b200_nonconv_t;

/* Poisson configuration of the projection step. The reference hard-codes CG with
 * poisson_solver_params_default() (solver_projection.c:217-218, linear_solver.c:37-47). */
typedef struct {
    poisson_solver_method_t method;  /* JACOBI, REDBLACK_SOR, CG, BICGSTAB */
    poisson_solver_params_t params;  /* tolerance, absolute_tolerance, max_iterations, omega,
                                        check_interval, preconditioner (CG only) */
    cfd_b200_nonconv_t on_nonconvergence;
} cfd_b200_poisson_config_t;

CFD_B200_API cfd_b200_poisson_config_t cfd_b200_poisson_config_default(void);
/* Process-wide configuration used by solve_projection_method_gpu and the factories.
 * Also settable with env CFD_B200_POISSON={jacobi,rbsor,cg,pcg,bicgstab,multigrid}. */
CFD_B200_API cfd_status_t cfd_b200_set_poisson_config(const cfd_b200_poisson_config_t* cfg);
CFD_B200_API cfd_b200_poisson_config_t cfd_b200_get_poisson_config(void);

/* Thread-local human-readable error of the last failing cfd_b200 / drop-in call. */
CFD_B200_API const char* cfd_b200_last_error(void);
CFD_B200_API const char* cfd_b200_version(void);
/* Number of kernels this library has launched since load (all streams, all sessions). */
CFD_B200_API unsigned long long cfd_b200_kernel_launch_count(void);

/* Pinned host memory helpers (for callers that want async H2D/D2H of their flow_field). */
CFD_B200_API void* cfd_b200_host_alloc(size_t bytes);
CFD_B200_API void cfd_b200_host_free(void* p);

/* ---- Device-resident session: one slab of the (possibly decomposed) grid on one GPU ---- */

typedef struct cfd_b200_session cfd_b200_session_t;

typedef struct {
    /* Local slab extents INCLUDING the one-cell boundary/halo layer on every side. */
    size_t nx, ny, nz;            /* nz == 1 for 2-D */
    double dx, dy, dz;            /* uniform spacing (dz ignored when nz == 1) */
    double xmin, ymin, zmin;      /* coordinate of local index 0 (source term uses x_i, y_j) */
    /* Slab decomposition along z (along y when nz == 1). rank 0 holds the low physical
     * boundary, rank nranks-1 the high one; interior ranks have halos on both sides. */
    int rank, nranks;
    int device;                   /* CUDA device ordinal, -1 = current */
} cfd_b200_session_desc_t;

CFD_B200_API cfd_b200_session_t* cfd_b200_session_create(const cfd_b200_session_desc_t* desc);
CFD_B200_API void cfd_b200_session_destroy(cfd_b200_session_t* s);

/* Multi-GPU: one process per GPU.  Rank 0 obtains a 128-byte ncclUniqueId, the caller
 * broadcasts it (e.g. torch.distributed) and every rank joins.  nccl_lib is the path of the
 * libnccl.so.2 to dlopen (NULL = default search).  Used for one-cell halo exchange
 * (ncclSend/ncclRecv) and the Krylov / residual all-reduces; all stream-ordered. */
CFD_B200_API cfd_status_t cfd_b200_nccl_unique_id(const char* nccl_lib, char out[128]);
CFD_B200_API cfd_status_t cfd_b200_session_init_comm(cfd_b200_session_t* s, const char* nccl_lib,
                                                     const char id[128]);

/* Host<->device movement of the local slab (host arrays sized nx*ny*nz, reference layout).
 * rho: only rho[0] is used by the projection; RK uses the full field. */
CFD_B200_API cfd_status_t cfd_b200_session_upload(cfd_b200_session_t* s, const flow_field* host);
CFD_B200_API cfd_status_t cfd_b200_session_download(cfd_b200_session_t* s, flow_field* host);
/* Same with a field mask: bit0 u, bit1 v, bit2 w, bit3 p, bit4 rho, bit5 T. */
CFD_B200_API cfd_status_t cfd_b200_session_upload_mask(cfd_b200_session_t* s, const flow_field* host, int mask);
CFD_B200_API cfd_status_t cfd_b200_session_download_mask(cfd_b200_session_t* s, flow_field* host, int mask);
/* Fill the device fields with the deterministic synthetic state of SURVEY 8(d) config C5
 * (sinusoidal velocity + splitmix64 perturbation); no host traffic. */
/* Cell coordinates x[nx], y[ny] of this slab for the default source term (A_u sin(pi y) ...,
 * cpu/solver_explicit_euler.c:326-331).  Default: xmin + i*dx, ymin + j*dy; a decomposed run
 * passes the slice of the GLOBAL grid arrays so every rank sees bit-identical coordinates. */
CFD_B200_API cfd_status_t cfd_b200_session_set_coords(cfd_b200_session_t* s, const double* x, const double* y);

CFD_B200_API cfd_status_t cfd_b200_session_init_synthetic(cfd_b200_session_t* s, uint64_t seed,
                                                          size_t global_k0, size_t global_nz);

/* Advance nsteps projection steps (solver_projection.c:111-290) fully on the device. */
CFD_B200_API cfd_status_t cfd_b200_session_projection(cfd_b200_session_t* s,
                                                      const ns_solver_params_t* params,
                                                      const cfd_b200_poisson_config_t* poisson,
                                                      int nsteps, ns_solver_stats_t* stats);
/* Advance nsteps explicit steps: order 1 = explicit Euler with the CPU path's limiters
 * (cpu/solver_explicit_euler.c:337-582), 2 = Heun, 4 = classical RK4
 * (cpu/solver_rk2.c, cpu/solver_rk4.c). */
CFD_B200_API cfd_status_t cfd_b200_session_rk(cfd_b200_session_t* s, int order,
                                              const ns_solver_params_t* params, int nsteps,
                                              ns_solver_stats_t* stats);
/* Stand-alone Poisson solve on device-resident x (the session's pressure) and rhs. */
CFD_B200_API cfd_status_t cfd_b200_session_poisson(cfd_b200_session_t* s,
                                                   const cfd_b200_poisson_config_t* poisson,
                                                   poisson_solver_stats_t* stats);
CFD_B200_API cfd_status_t cfd_b200_session_upload_poisson(cfd_b200_session_t* s, const double* x,
                                                          const double* rhs);
CFD_B200_API cfd_status_t cfd_b200_session_download_poisson(cfd_b200_session_t* s, double* x);

/* Poisson statistics of the most recent projection step / Poisson solve. */
CFD_B200_API poisson_solver_stats_t cfd_b200_session_last_poisson_stats(const cfd_b200_session_t* s);

/* ---- Multi-GPU behind the drop-in boundary (SURVEY 8e "process model") ----
 * The reference's entry points receive one host flow_field (gpu_device.h:209-249, called from
 * solver_registry.c:1121-1245 behind run_simulation_step).  With a device count G > 1 -- this setter, or the
 * environment variable CFD_B200_NGPU when the setter was never called (0 = back to the environment) -- every
 * solve_*_method_gpu call and every ns_solver_t made by the cfd_b200_create_*_gpu_solver factories cuts the grid
 * into G slabs along z (along y when nz == 1; uneven when the interior layers do not divide), one GPU and one
 * host thread per slab inside this process, NCCL communicators from ncclCommInitAll.  Clamped to the visible
 * devices and to >= 4 owned layers per slab; the multigrid method stays on one GPU. */
CFD_B200_API cfd_status_t cfd_b200_set_device_count(int ngpu);
CFD_B200_API int cfd_b200_get_device_count(void);
/* Position of a session's slab in the global grid: global layer count of the decomposed axis (boundary
 * layers included) and the global index of local layer 0.  Needed for uneven slabs (red/black colouring,
 * optimal omega); uniform slabs derive both from rank / nranks. */
CFD_B200_API cfd_status_t cfd_b200_session_set_slab(cfd_b200_session_t* s, size_t global_layers, size_t global_lo);
/* Device buffers of the stateless solve_*_method_gpu symbols are cached between calls (LRU of
 * CFD_B200_SESSION_CACHE = 2 shapes); this frees the idle ones. */
CFD_B200_API void cfd_b200_release_cached_sessions(void);

/* ---- Arithmetic policy of the stencil and Poisson kernels (csrc/common.cuh) ----
 * exact (0, default): no FMA contraction, the reference's divisions reproduced -- Jacobi / Red-Black SOR / RK /
 * Euler / predictor / corrector results are bit-identical to the reference's CPU path.
 * fast (1): FMA contraction and reciprocal multiplies instead of divisions -- same formulas, results within
 * ~1e-13 per step of the exact build (tests/test_gpu_fast.py holds it to the north-star tolerances), fewer FP64
 * instructions in the issue-bound kernels.  -1 = decide from the environment (CFD_B200_ARITH=fast|exact).
 * Applies to sessions / solver objects created afterwards. */
CFD_B200_API cfd_status_t cfd_b200_set_arithmetic(int fast);
CFD_B200_API int cfd_b200_session_arithmetic(const cfd_b200_session_t* s);

/* ---- Peer-memory transport between slabs (csrc/peer.cu) ----
 * Halo layers and the <= 8-word all-reduces of the Krylov / residual scalars travel by direct stores into the
 * neighbours' device memory over NVLink, with sequence flags instead of a communication library.  Slabs of one
 * process (cfd_b200_set_device_count) are connected automatically; one process per GPU: every rank exports a
 * 64-byte handle, the caller gathers them in rank order (nranks x 64 bytes) and every rank attaches.  After a
 * successful attach the session uses this transport; otherwise the NCCL communicator of
 * cfd_b200_session_init_comm.  cfd_b200_session_transport: 0 none, 1 NCCL, 2 peer memory. */
CFD_B200_API cfd_status_t cfd_b200_session_peer_export(cfd_b200_session_t* s, char out[64]);
CFD_B200_API cfd_status_t cfd_b200_session_peer_attach(cfd_b200_session_t* s, const char* handles);
CFD_B200_API int cfd_b200_session_transport(const cfd_b200_session_t* s);

/* ---- Host-sync policy of an ns_solver_t made by the cfd_b200_create_*_gpu_solver factories ----
 * FULL (default): every step()/solve() uploads the caller's fields and downloads the result -- a drop-in with
 * no assumptions.  UPLOAD_BOUNDARY: the caller promises that, since the previous call on this solver returned,
 * it has edited only boundary cells (the outermost layer) of the SAME field -- what the reference's drivers do
 * (tests/validation/lid_driven_cavity_common.h:306-338); the first call uploads everything, later calls only
 * the six faces.  DOWNLOAD_LAZY: the result stays on the device (boundary velocities set by the caller are
 * preserved by the steppers, boundary_copy_utils.h:93-148) and nothing is uploaded while the host copy is the
 * stale one; cfd_b200_solver_sync_host() brings the caller's arrays up to date. */
typedef enum {
    CFD_B200_SYNC_FULL = 0,
    CFD_B200_SYNC_UPLOAD_BOUNDARY = 1,
    CFD_B200_SYNC_DOWNLOAD_LAZY = 2
} cfd_b200_host_sync_t;
CFD_B200_API cfd_status_t cfd_b200_solver_set_host_sync(ns_solver_t* solver, int flags);
CFD_B200_API cfd_status_t cfd_b200_solver_sync_host(ns_solver_t* solver, flow_field* field);

/* ---- An entry point that allocates nothing (csrc/no_alloc.cu) ----
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * L-infinity residual max |lap x - rhs| over the interior cells of a uniform grid -- the quantity of the reference's
 * poisson_solver_compute_residual (linear_solver.c:304-346), same expression and bits as the initial / final
 * residual the stationary GPU solvers report.  d_x, d_rhs: DEVICE arrays of nx*ny*nz doubles in the reference layout
 * (boundary cells included; nz == 1 for 2-D, dz ignored then); d_out: DEVICE pointer to one double, written when the
 * work queued on `stream` has run (the call does not synchronise).  The per-block partial maxima go through a
 * file-scope static device array of 524 544 bytes that every GPU loading the library holds until exit: grids of more
 * than CFD_B200_NOALLOC_MAX_CELLS cells are rejected with CFD_ERROR_LIMIT_EXCEEDED (use a session for those).  The
 * call makes no cudaMalloc / cudaFree and no allocating library call.  All calls on one GPU share the scratch:
 * calls that OVERLAP on the GPU (different streams without ordering) race and may return garbage. */
#define CFD_B200_NOALLOC_MAX_CELLS ((size_t)1 << 24) /* 16 777 216 cells: 256^3, 4096^2 */
CFD_B200_API size_t cfd_b200_noalloc_max_cells(void);
CFD_B200_API cfd_status_t cfd_b200_poisson_residual_linf_noalloc(const double* d_x, const double* d_rhs, size_t nx,
                                                                 size_t ny, size_t nz, double dx, double dy, double dz,
                                                                 double* d_out, cfd_b200_stream_t stream);

/* ---- Measurement hooks (bench.py): per-kernel CUDA-event timing on the session stream ---- */
#define CFD_B200_MAX_KERNEL_KINDS 48
typedef struct {
    const char* name;            /* kernel family name */
    unsigned long long launches; /* launches since reset */
    double total_ms;             /* summed CUDA-event time (only when profiling enabled) */
    double bytes_per_cell;       /* algorithmic bytes per interior cell per launch */
} cfd_b200_kernel_stat_t;
CFD_B200_API void cfd_b200_session_profile(cfd_b200_session_t* s, int enable);
CFD_B200_API int cfd_b200_session_kernel_stats(const cfd_b200_session_t* s,
                                               cfd_b200_kernel_stat_t* out, int max);
CFD_B200_API void cfd_b200_session_reset_kernel_stats(cfd_b200_session_t* s);
/* cudaStream_t of the session as void* (so callers can record their own events on it). */
CFD_B200_API void* cfd_b200_session_stream(cfd_b200_session_t* s);
CFD_B200_API cfd_status_t cfd_b200_session_sync(cfd_b200_session_t* s);

#ifdef __cplusplus
}
#endif
#endif /* CFD_B200_H */
