/* solo12_model.h -- rigid-body constants of Solo12 used by the whole-body kernels and by the oracle.
 *
 * The reference loads these through Pinocchio from example-robot-data's solo12.urdf
 * (/root/reference/scripts/QP_WBC.py:20-21, scripts/solo12InvKin.py:12-13); neither is in the reference tree nor in
 * this image, so the numbers below are the builder's transcription of the public URDF, pinned to the
 * anchors the reference tree does contain:
 *   hip (HAA) positions (+-0.1946, +-0.0875), lateral offsets 0.014 + 0.03745 + 0.008 => shoulder-to-foot lateral
 *   0.14695 (scripts/Controller.py:131-133), leg segments 0.16 + 0.16 (scripts/test_mpc.py:39-40),
 *   total mass 2.50000279 (src/MPC.cpp:17) -- the lower-leg mass (with foot) is set so that the total matches --
 *   and the base-link inertia diag (0.00578574, 0.01938108, 0.02476124) (scripts/crocoddyl_class/MPC_crocoddyl.py:32).
 * PARITY UNPINNED: results that depend on these constants cannot be checked against the reference here.
 *
 * Legs in the reference's order FL, FR, HL, HR; sx = +1 front / -1 hind, sy = +1 left / -1 right.
 * Joints per leg: HAA (axis x), HFE (axis y), KFE (axis y); then a fixed foot frame.
 */
#ifndef SOLO12_MODEL_H
#define SOLO12_MODEL_H

#define S12_GRAVITY 9.81

#define S12_BASE_MASS 1.16115091
#define S12_BASE_IXX 0.00578574
#define S12_BASE_IYY 0.01938108
#define S12_BASE_IZZ 0.02476124

/* joint placements in the parent link frame (multiply x by sx, y by sy) */
#define S12_HAA_X 0.1946
#define S12_HAA_Y 0.0875
#define S12_HFE_Y 0.014
#define S12_KFE_Y 0.03745
#define S12_KFE_Z (-0.16)
#define S12_FOOT_Y 0.008
#define S12_FOOT_Z (-0.16)

/* link 0: hip assembly (moves with HAA); com x mirrors with -sx, com y with sy; ixy mirrors with -sx*sy */
#define S12_L0_MASS 0.14853845
#define S12_L0_CX (-0.078707)
#define S12_L0_CY 0.01
#define S12_L0_CZ 0.0
#define S12_L0_IXX 0.00003024
#define S12_L0_IXY 0.00004671
#define S12_L0_IXZ 0.0
#define S12_L0_IYY 0.00041193
#define S12_L0_IYZ 0.0
#define S12_L0_IZZ 0.00041107
/* link 1: upper leg (moves with HFE); com y mirrors with sy; iyz mirrors with sy */
#define S12_L1_MASS 0.14737324
#define S12_L1_CX 0.0000153
#define S12_L1_CY 0.0176764
#define S12_L1_CZ (-0.0783823)
#define S12_L1_IXX 0.0004154
#define S12_L1_IXY 0.0
#define S12_L1_IXZ 0.0000001
#define S12_L1_IYY 0.00041637
#define S12_L1_IYZ (-0.00004589)
#define S12_L1_IZZ 0.00002982
/* link 2: lower leg + foot (moves with KFE) */
#define S12_L2_MASS 0.03880128
#define S12_L2_CX 0.0
#define S12_L2_CY 0.00776716
#define S12_L2_CZ (-0.07003876)
#define S12_L2_IXX 0.00008508
#define S12_L2_IXY 0.0
#define S12_L2_IXZ 0.0
#define S12_L2_IYY 0.0000858
#define S12_L2_IYZ (-0.000002)
#define S12_L2_IZZ 0.00000139

#endif
