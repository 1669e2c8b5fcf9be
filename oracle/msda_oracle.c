/*
 * oracle/msda_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU oracle for the MSDeformAttn hot path: a plain-C restatement of the reference's
 * ms_deform_attn_core_pytorch (pdvc/ops/functions/ms_deform_attn_func.py:41-68) and of
 * its autograd backward.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker / the timed CPU
 * baseline -- never as the product path (the product path is the CUDA library behind
 * include/msda.h and fails loudly when that library is missing).
 *
 * Pinning: the reference stores no golden vectors for this path (pdvc/ops/test.py prints,
 * never asserts, and is CUDA-only).  This oracle is pinned against outputs of the reference
 * function itself, imported unmodified from /root/reference and run in fp64 by
 * tests/golden/make_golden.py; the resulting fixtures are committed under tests/golden/ and
 * tests/test_oracle.py checks this file against them (<= 1e-12 relative).
 *
 * Built by oracle/Makefile into oracle/_build/libmsda_oracle.so (gcc -O2 -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define REAL double
#define FN(name) CAT(name, _f64)
#define FLOOR(x) floor(x)
#include "msda_oracle_body.inc"
#undef REAL
#undef FN
#undef FLOOR

#define REAL float
#define FN(name) CAT(name, _f32)
#define FLOOR(x) floorf(x)
#include "msda_oracle_body.inc"
#undef REAL
#undef FN
#undef FLOOR

int msda_oracle_version(void) { return 1; }
