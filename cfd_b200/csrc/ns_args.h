// ns_args.h -- small by-value argument bundles of the RK stage kernel.
#pragma once
namespace cfdb {
struct P4 {
    const double* a[4];
};
struct W4 {
    double* a[4];
};
struct RkArgs {
    double fac_next;   // Q_{s+1} = Q0 + fac_next * k           (not last stage)
    double w_k;        // acc += w_k * k  (first stage initialises acc instead)
    double w_final;    // Q = Q0 + w_final * (acc + k)         (last stage)
    int first, last;
    int wrap_y, wrap_z;  // wrap neighbour indices locally (axis not decomposed over ranks)
    double src_decay;    // exp(-lambda * iter * dt)
};
struct EulerArgs {
    double dt;         // min(params->dt, DT_CONSERVATIVE_LIMIT = 1e-4)  (solver_explicit_euler.c:395)
    double p_factor;   // -PRESSURE_UPDATE_FACTOR * dt, the leading product of the pressure increment (:518)
    double src_decay;  // exp(-lambda * iter * dt)
};
}  // namespace cfdb
