// jaco_model.h -- fp32 model constants of the compiled Jaco scene, as the kernels see them.
// Filled on the host from the flat float64 blob produced by mujoco_jaco_b200/model/compiler.py
// (layout: include/jaco_model_layout.h) and copied once into __constant__ memory.
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/jaco_model_layout.h"

#define JACO_NQ 23
#define JACO_NV 21
#define JACO_NU 9
#define JACO_NOBS 26
#define JACO_NAUX 72
#define JACO_NSTATE (JACO_NQ + JACO_NV + JACO_NAUX)  // rows of the SoA state buffer
#define JACO_MAXCON 32            /* cooperative path: lanes map to contacts */
#define JACO_MAXCON_GENERAL 48    /* general thread-per-env path (resets with the hand spawned inside / closed around the object: 34-36 contacts) */
#define JACO_NCH 9
#define JACO_NBODY 12  // 0 world, 1..9 chain, 10 object, 11 dest

enum { JACO_TASK_REACHING = 0, JACO_TASK_GRASPING = 1, JACO_TASK_PICKING = 2, JACO_TASK_PLACING = 3, JACO_TASK_PICKANDPLACE = 4, JACO_TASK_RELEASING = 5, JACO_TASK_CARRYING = 6 };
enum { JACO_GEOM_PLANE = 0, JACO_GEOM_BOX = 1, JACO_GEOM_CAPSULE = 2 };

// aux[] slots of the per-env state (shared with oracle/jaco_oracle.c jo_get_state)
enum {
  AUX_TARGET = 0, AUX_GRIP = 6, AUX_ACTION_IN = 7, AUX_PICKED = 8, AUX_CUR_STEPS = 9, AUX_NUM_EPISODES = 10,
  AUX_OBJ_GOAL = 11, AUX_DEST_GOAL = 14, AUX_REACH_GOAL = 17, AUX_RNG = 23, AUX_TOUCH = 24, AUX_WARM = 25,
  AUX_LAGQ = 46, AUX_LAGV = 55, AUX_ACCUM_REW = 64, AUX_EPLEN = 65,
  AUX_RNG_HI = 66   // Philox step counter = AUX_RNG (low 24 bits) + 2^24 * AUX_RNG_HI: exact in float rows for all uint32
};

struct JacoGeom { int type, body; float pos[3], R[9], size[3], rbound; };
struct JacoPair { int g1, g2, group, condim; float friction[3], solref[2], solimp[5], margin, invw_tran, invw_rot; };
struct JacoSite { int type, body, geom; float pos[3], R[9], size[3]; };

struct JacoModel {
  float ch_pos[9][3], ch_quat[9][4], ch_axis[9][3], ch_ref[9], ch_range[9][2], ch_damping[9];
  int ch_limited[9];
  float mass[JACO_NBODY], com[JACO_NBODY][3], inertia[JACO_NBODY][6];  // body frame, about COM: xx yy zz xy xz yz
  float free_qpos0[2][7];
  float motor_fr[6][2], finger_act[3][5];
  float ee_off[3], ee_obj_off[3], base_pos[3], base_link_pos[3], base_link_eul[3];
  float dof_invw[JACO_NV], meaninertia, hand_sphere[4];
  float body_sphere[4][4];  // bounding spheres (body frame: centre xyz, radius) of the collision geoms on bodies 6..9
  float dt, gravz;
  unsigned long long pm_boxbox, pm_capbox, pm_plane;  // arm pairs (index >= 4) by narrow-phase routine: bit p set
  JacoGeom geoms[JM_NGEOM];
  JacoPair pairs[JM_NPAIR];
  JacoSite sites[JM_NSITE];
};

static inline void jaco_model_from_blob(const double* b, JacoModel* m) {
  for (int k = 0; k < 9; k++) {
    for (int i = 0; i < 3; i++) {
      m->ch_pos[k][i] = (float)b[JM_CHAIN_POS + 3 * k + i];
      m->ch_axis[k][i] = (float)b[JM_CHAIN_AXIS + 3 * k + i];
      m->com[k + 1][i] = (float)b[JM_CHAIN_COM + 3 * k + i];
    }
    for (int i = 0; i < 4; i++) m->ch_quat[k][i] = (float)b[JM_CHAIN_QUAT + 4 * k + i];
    m->ch_ref[k] = (float)b[JM_CHAIN_REF + k];
    m->ch_range[k][0] = (float)b[JM_CHAIN_RANGE + 2 * k];
    m->ch_range[k][1] = (float)b[JM_CHAIN_RANGE + 2 * k + 1];
    m->ch_limited[k] = b[JM_CHAIN_LIMITED + k] != 0;
    m->ch_damping[k] = (float)b[JM_CHAIN_DAMPING + k];
    m->mass[k + 1] = (float)b[JM_CHAIN_MASS + k];
    for (int i = 0; i < 6; i++) m->inertia[k + 1][i] = (float)b[JM_CHAIN_INERTIA + 6 * k + i];
  }
  m->mass[0] = 0;
  for (int i = 0; i < 3; i++) m->com[0][i] = 0;
  for (int i = 0; i < 6; i++) m->inertia[0][i] = 0;
  for (int f = 0; f < 2; f++) {
    m->mass[10 + f] = (float)b[JM_FREE_MASS + f];
    for (int i = 0; i < 3; i++) m->com[10 + f][i] = (float)b[JM_FREE_COM + 3 * f + i];
    for (int i = 0; i < 6; i++) m->inertia[10 + f][i] = (float)b[JM_FREE_INERTIA + 6 * f + i];
    for (int i = 0; i < 7; i++) m->free_qpos0[f][i] = (float)b[JM_FREE_QPOS0 + 7 * f + i];
  }
  for (int i = 0; i < 6; i++) { m->motor_fr[i][0] = (float)b[JM_MOTOR_FORCERANGE + 2 * i]; m->motor_fr[i][1] = (float)b[JM_MOTOR_FORCERANGE + 2 * i + 1]; }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 5; j++) m->finger_act[i][j] = (float)b[JM_FINGER_ACT + 5 * i + j];
  for (int i = 0; i < 3; i++) {
    m->ee_off[i] = (float)b[JM_EE_OFFSET + i]; m->ee_obj_off[i] = (float)b[JM_EE_OBJ_OFFSET + i];
    m->base_pos[i] = (float)b[JM_BASE_POSITION + i]; m->base_link_pos[i] = (float)b[JM_BASE_LINK_POSITION + i]; m->base_link_eul[i] = (float)b[JM_BASE_LINK_EULER + i];
  }
  for (int i = 0; i < JACO_NV; i++) m->dof_invw[i] = (float)b[JM_DOF_INVWEIGHT0 + i];
  m->meaninertia = (float)b[JM_MEANINERTIA];
  for (int i = 0; i < 4; i++) m->hand_sphere[i] = (float)b[JM_HAND_SPHERE + i];
  m->dt = (float)b[JM_OPT]; m->gravz = (float)b[JM_OPT + 1];
  for (int g = 0; g < JM_NGEOM; g++) {
    const double* r = b + JM_GEOMS + JM_NGEOM_FIELDS * g; JacoGeom* G = &m->geoms[g];
    G->type = (int)r[0]; G->body = (int)r[1];
    for (int i = 0; i < 3; i++) { G->pos[i] = (float)r[2 + i]; G->size[i] = (float)r[14 + i]; }
    for (int i = 0; i < 9; i++) G->R[i] = (float)r[5 + i];
    G->rbound = (float)r[17];
  }
  for (int p = 0; p < JM_NPAIR; p++) {
    const double* r = b + JM_PAIRS + JM_NPAIR_FIELDS * p; JacoPair* P = &m->pairs[p];
    P->g1 = (int)r[0]; P->g2 = (int)r[1]; P->group = (int)r[2]; P->condim = (int)r[3];
    for (int i = 0; i < 3; i++) P->friction[i] = (float)r[4 + i];
    P->solref[0] = (float)r[7]; P->solref[1] = (float)r[8];
    for (int i = 0; i < 5; i++) P->solimp[i] = (float)r[9 + i];
    P->margin = (float)r[14]; P->invw_tran = (float)r[15]; P->invw_rot = (float)r[16];
  }
  for (int s = 0; s < JM_NSITE; s++) {
    const double* r = b + JM_SITES + JM_NSITE_FIELDS * s; JacoSite* S = &m->sites[s];
    S->type = (int)r[0]; S->body = (int)r[1]; S->geom = (int)r[17];
    for (int i = 0; i < 3; i++) { S->pos[i] = (float)r[2 + i]; S->size[i] = (float)r[14 + i]; }
    for (int i = 0; i < 9; i++) S->R[i] = (float)r[5 + i];
  }
  m->pm_boxbox = m->pm_capbox = m->pm_plane = 0;
  for (int p = 4; p < JM_NPAIR; p++) {
    const int t1 = m->geoms[m->pairs[p].g1].type, t2 = m->geoms[m->pairs[p].g2].type;
    const unsigned long long bit = 1ull << p;
    if (t1 == JACO_GEOM_PLANE || t2 == JACO_GEOM_PLANE) m->pm_plane |= bit;
    else if (t1 == JACO_GEOM_BOX && t2 == JACO_GEOM_BOX) m->pm_boxbox |= bit;
    else m->pm_capbox |= bit;
  }
  // per-body bounding spheres of the hand cluster (hand + three fingers), used by the fast path's proximity tests
  for (int b = 6; b <= 9; b++) {
    double c[3] = {0, 0, 0}; int cnt = 0;
    for (int g = 0; g < JM_NGEOM; g++) if (m->geoms[g].body == b) { for (int i = 0; i < 3; i++) c[i] += m->geoms[g].pos[i]; cnt++; }
    for (int i = 0; i < 3; i++) c[i] /= (cnt ? cnt : 1);
    double rad = 0;
    for (int itn = 0; itn < 400; itn++) {
      int far = -1; rad = 0;
      for (int g = 0; g < JM_NGEOM; g++) if (m->geoms[g].body == b) {
        double d = 0; for (int i = 0; i < 3; i++) d += (m->geoms[g].pos[i] - c[i]) * (m->geoms[g].pos[i] - c[i]);
        d = sqrt(d) + m->geoms[g].rbound;
        if (d > rad) { rad = d; far = g; }
      }
      if (far < 0) break;
      double step = 0.5 / (itn + 2);
      for (int i = 0; i < 3; i++) c[i] += step * (m->geoms[far].pos[i] - c[i]);
    }
    rad = 0;
    for (int g = 0; g < JM_NGEOM; g++) if (m->geoms[g].body == b) {
      double d = 0; for (int i = 0; i < 3; i++) d += (m->geoms[g].pos[i] - c[i]) * (m->geoms[g].pos[i] - c[i]);
      d = sqrt(d) + m->geoms[g].rbound; if (d > rad) rad = d;
    }
    for (int i = 0; i < 3; i++) m->body_sphere[b - 6][i] = (float)c[i];
    m->body_sphere[b - 6][3] = (float)(rad * 1.0001 + 1e-6);
  }
}
