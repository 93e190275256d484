/*
 * abi_probe.c -- compiled INTO oracle/_ref/libcfd_ref*.so together with the real reference
 * headers.  Exports the sizeof/offsetof table of every reference type that crosses the
 * drop-in boundary, in the same order as cfd_b200_abi_table() in the product library
 * (cfd_b200/csrc/abi_table.inc), so tests can prove include/cfd_b200.h mirrors the reference.
 * TEST INFRASTRUCTURE ONLY.
 */
#include "cfd/core/gpu_device.h"
#include "cfd/solvers/navier_stokes_solver.h"
#include "cfd/solvers/poisson_solver.h"
#include <stddef.h>

#define CFD_B200_ABI_FN ref_abi_table
#include "../cfd_b200/csrc/abi_table.inc"
