// Physics constants of the CUDA path. The reference never overrides pybullet's defaults (no setTimeStep /
// setPhysicsEngineParameter under /root/reference/urgym), so these are Bullet 3.2.x / pybullet defaults as recalled
// (pybullet is not installable here: parity with pybullet itself is UNPINNED) plus the values the
// reference passes explicitly (robot.py:211 gear maxForce/erp; env_ball_balance_v0.py:94-103 settle loop).
#pragma once
#include "real.h"

#define UGP_DT ((real)(1.0 / 240.0))
#define UGP_INV_DT 240.0_r
#define UGP_SOLVER_ITERS 50
#define UGP_RESIDUAL_THRESHOLD 1e-7_r
#define UGP_ERP 0.2_r
#define UGP_ERP2 0.08_r
#define UGP_LINEAR_SLOP 1e-5_r
#define UGP_RESTITUTION_VEL_THRESHOLD 0.2_r
#define UGP_MAX_FRICTION 10.0_r
#define UGP_MAX_COORD_VEL 100.0_r
#define UGP_LIN_DAMPING 0.04_r
#define UGP_ANG_DAMPING 0.04_r
#define UGP_CONTACT_BREAKING 0.02
#define UGP_ANGULAR_MOTION_THRESHOLD (0.5_r * 1.5707963267948966_r)
#define UGP_MOTOR_KP 0.1_r
#define UGP_MOTOR_KD 1.0_r
#define UGP_LIMIT_MAX_IMPULSE 100.0_r
#define UGP_GEAR_MAX_FORCE 100.0_r
#define UGP_GEAR_ERP 1.0_r
#define UGP_IK_ITERS 20
#define UGP_IK_DAMPING 0.5_r
#define UGP_IK_RESIDUAL 1e-4_r
#define UGP_IK_MAX_STEP 0.7853981633974483_r
#define UGP_SETTLE_MAX_TICKS 480
#define UGP_SETTLE_THRESHOLD 5e-3_r
#define UGP_RESET_ARM_TICKS 10
#define UGP_CLUB_ZC ((-0.09 + 0.08) / 3.0)
#define UGP_PI 3.14159265358979323846_r
// contact-free ticks with this many joint rows, or one more, run the register-resident sweeps (0 disables them):
// 7 motors + 5 gears = 12
#ifndef UG_JREG
#define UG_JREG 12
#endif
// resident CTAs per SM asked of ptxas: 13 fp32 envs fit the shared memory of an SM; the parity build takes the registers it wants
#ifdef UG_F64
#define UG_MIN_CTAS 3
#else
#define UG_MIN_CTAS 13
#endif
