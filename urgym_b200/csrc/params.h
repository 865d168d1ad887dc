// Physics constants of the CUDA path. The reference never overrides pybullet's defaults (no setTimeStep /
// setPhysicsEngineParameter under /root/reference/urgym), so these are Bullet 3.2.x / pybullet defaults as recalled
// (pybullet is not installable here: parity with pybullet itself is UNPINNED) plus the values the
// reference passes explicitly (robot.py:211 gear maxForce/erp; env_ball_balance_v0.py:94-103 settle loop).
#pragma once

#define UGP_DT (1.0f / 240.0f)
#define UGP_INV_DT 240.0f
#define UGP_SOLVER_ITERS 50
#define UGP_RESIDUAL_THRESHOLD 1e-7f
#define UGP_ERP 0.2f
#define UGP_ERP2 0.08f
#define UGP_LINEAR_SLOP 1e-5f
#define UGP_RESTITUTION_VEL_THRESHOLD 0.2f
#define UGP_MAX_FRICTION 10.0f
#define UGP_MAX_COORD_VEL 100.0f
#define UGP_LIN_DAMPING 0.04f
#define UGP_ANG_DAMPING 0.04f
#define UGP_CONTACT_BREAKING 0.02
#define UGP_ANGULAR_MOTION_THRESHOLD (0.5f * 1.5707963267948966f)
#define UGP_MOTOR_KP 0.1f
#define UGP_MOTOR_KD 1.0f
#define UGP_LIMIT_MAX_IMPULSE 100.0f
#define UGP_GEAR_MAX_FORCE 100.0f
#define UGP_GEAR_ERP 1.0f
#define UGP_IK_ITERS 20
#define UGP_IK_DAMPING 0.5f
#define UGP_IK_RESIDUAL 1e-4f
#define UGP_IK_MAX_STEP 0.7853981633974483f
// capacity choice shared with the oracle (bullet_params.h): rolling-friction rows for the first 3 rolling manifolds of a tick
#define UGP_MAX_ROLLING_CONTACTS 3
#define UGP_SETTLE_MAX_TICKS 480
#define UGP_SETTLE_THRESHOLD 5e-3f
#define UGP_RESET_ARM_TICKS 10
#define UGP_CLUB_ZC ((-0.09 + 0.08) / 3.0)
#define UGP_PI 3.14159265358979323846f
// contact-free ticks with this many joint rows, or one more, run the register-resident sweeps (0 disables them):
// 7 motors + 5 gears = 12
#ifndef UG_JREG
#define UG_JREG 12
#endif
