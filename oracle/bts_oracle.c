/*
 * bts_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE. See bts_oracle_impl.h.
 * Instantiates the CPU restatement of the reference hot path in fp32 and fp64.
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off: no FMA contraction, so the fp32 LPG forward
 * reproduces the reference's separately rounded mul/add sequence bit for bit).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdlib.h>
#include <stddef.h>

#define REAL float
#define SUF f32
#define EXP expf
#define SIN sinf
#define COS cosf
#define SQRT sqrtf
#define FABS fabsf
#include "bts_oracle_impl.h"
#undef REAL
#undef SUF
#undef EXP
#undef SIN
#undef COS
#undef SQRT
#undef FABS

#define REAL double
#define SUF f64
#define EXP exp
#define SIN sin
#define COS cos
#define SQRT sqrt
#define FABS fabs
#include "bts_oracle_impl.h"
