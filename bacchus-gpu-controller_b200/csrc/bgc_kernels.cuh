// bgc_kernels.cuh — sm_100a device code of libbgc.so (umbrella header).
//
// K0 policy_eval  : batched form of mutate(), reference src/admission.rs:241-431   (row A1)  bgc_k0.cuh
// K1 usage_reduce : per-namespace wrapping-i64 usage, no reference function        (row A3)  bgc_k1.cuh
// K2 quota_eval   : snapshot quota verdict per candidate pod, no reference function (row A4)  bgc_k2.cuh
// K3 admit_greedy : sequential in-batch admission, no reference function           (row N1)  bgc_k3.cuh
// K1/K2 over narrow (typed) columns: the same kernels with widening loads                  bgc_typed.cuh
// plus the counter-based synthetic generator and stream checksums                            bgc_synth_kernels.cuh
//
// All of them are HBM-streaming integer kernels: no tensor cores, no TMEM.  What matters is
// whole-sector 128/256-bit coalesced read-only loads that bypass L1, enough loads in flight per
// thread, grids sized for the 148 SMs, and - for K1 - keeping the atomic traffic per pod low.
#pragma once
#include "bgc_common.cuh"
#include "bgc_k0.cuh"
#include "bgc_k1.cuh"
#include "bgc_k2.cuh"
#include "bgc_k3.cuh"
#include "bgc_typed.cuh"
#include "bgc_synth_kernels.cuh"
