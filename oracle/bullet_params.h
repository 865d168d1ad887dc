// TEST INFRASTRUCTURE — part of the CPU oracle. Not linked into, imported by, or executed from the product path.
//
// PARITY UNPINNED: pybullet (the third-party module that owns all of the reference's arithmetic, required unpinned
// at /root/reference/setup.py:9) is not installable in the build container or on the GPU box, and the reference
// holds no tests or golden vectors. Every constant below marked [recalled] is the author's recollection of
// Bullet 3.2.x / pybullet defaults that the reference inherits by never overriding them
// (no setTimeStep / setPhysicsEngineParameter anywhere under /root/reference/urgym). They are gathered in this one
// header so that they can be corrected in one place on a machine that has pybullet.
#pragma once

// --- world defaults (btContactSolverInfo as configured by pybullet's PhysicsServerCommandProcessor) [recalled]
#define BP_DT (1.0 / 240.0)               // fixed time step, no sub-steps
#define BP_SOLVER_ITERS 50                // m_numIterations
#define BP_RESIDUAL_THRESHOLD 1e-7        // m_leastSquaresResidualThreshold (max over rows of squared velocity change)
#define BP_ERP 0.2                        // m_erp: joint-limit / gear positional error reduction
#define BP_ERP2 0.08                      // m_erp2: contact penetration error reduction ("contactERP")
#define BP_LINEAR_SLOP 1e-5               // m_linearSlop
#define BP_RESTITUTION_VEL_THRESHOLD 0.2  // m_restitutionVelocityThreshold
#define BP_MAX_FRICTION 10.0              // btManifoldResult MAX_FRICTION clamp on combined coefficients
#define BP_MAX_COORD_VEL 100.0            // btMultiBody::m_maxCoordinateVelocity
#define BP_LIN_DAMPING 0.04               // btMultiBody::m_linearDamping  (force = m v (k + k |v|))
#define BP_ANG_DAMPING 0.04               // btMultiBody::m_angularDamping (torque = I w (k + k |w|))
#define BP_CONTACT_BREAKING 0.02          // gContactBreakingThreshold, scaled by the shape's angular motion disc
#define BP_ANGULAR_MOTION_THRESHOLD (0.5 * 1.5707963267948966)

// --- joint motor defaults of setJointMotorControl2(POSITION_CONTROL) [recalled]; call sites robot.py:127-128,217-218
#define BP_MOTOR_KP 0.1
#define BP_MOTOR_KD 1.0
// --- joint limit constraint (btMultiBodyJointLimitConstraint) [recalled]
#define BP_LIMIT_MAX_IMPULSE 100.0
// --- gear constraint parameters given by the reference: robot.py:211 (maxForce=100, erp=1)
#define BP_GEAR_MAX_FORCE 100.0
#define BP_GEAR_ERP 1.0
// --- IK (pybullet calculateInverseKinematics -> BussIK DLS with orientation) [recalled]; call site robot.py:119-121
#define BP_IK_ITERS 20                    // maxNumIterations passed by the reference
#define BP_IK_DAMPING 0.5                 // default per-joint damping added to the diagonal of J^T J
#define BP_IK_RESIDUAL 1e-4               // default residualThreshold (position error)
#define BP_IK_MAX_STEP 0.7853981633974483 // BussIK MaxAngleDLS = 45 degrees
// --- settle loop given by the reference: env_ball_balance_v0.py:94-103
#define BP_SETTLE_MAX_TICKS 480
#define BP_SETTLE_THRESHOLD 5e-3
#define BP_RESET_ARM_TICKS 10             // robot.py:97-98

// --- run-time copies of the recalled values, so that tools/ablate_params.py can ask how sensitive the replay of the
//     reference's pybullet-trained actors is to each of them. Defaults are the macros above; nothing else
//     changes them.
struct BulletParams {
  double erp = BP_ERP, erp2 = BP_ERP2, linear_slop = BP_LINEAR_SLOP, restitution_threshold = BP_RESTITUTION_VEL_THRESHOLD;
  double lin_damping = BP_LIN_DAMPING, ang_damping = BP_ANG_DAMPING, contact_breaking = BP_CONTACT_BREAKING;
  double motor_kp = BP_MOTOR_KP, motor_kd = BP_MOTOR_KD, ik_damping = BP_IK_DAMPING, max_friction = BP_MAX_FRICTION;
  double solver_iters = BP_SOLVER_ITERS;
  double robot_mu = 0.5;        // URDF links without <contact> coefficients [recalled default 0.5]
  double body_mu = 0.5;         // createMultiBody bodies (paddle, balls, table, pitch, club parts) [recalled default 0.5]
  double cube_mu = 1.0;         // cube.urdf lateral_friction [recalled]
  double plane_mu = 1.0;        // plane.urdf lateral_friction [recalled]
  double cube_mass = 1.0;       // cube.urdf mass under globalScaling=0.04 [recalled: mass not scaled]
  double friction_cone = 1.0;   // 1: implicit cone over the two tangent rows; 0: independent box limits per row
  double roll_points = 1.0;     // rolling-friction rows on the first `roll_points` points of a manifold
  double hull_links = 1.0;      // 1: gripper links with mesh geometry collide as convex polytopes; 0: as their fitted boxes (round 1)
  double disc_half = 0.04;      // ablation: half side of the box standing in for the club's top disc
  double link_mask = 524287.0;  // ablation: bit i = robot link i collides (all 19 by default)
  double box_shrink = 0.0;      // ablation: metres taken off every half extent of the fitted robot link boxes
};
inline BulletParams& bp() { static BulletParams p; return p; }
