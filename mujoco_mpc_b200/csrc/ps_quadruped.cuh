// Quadruped Flat residual device function (mjpc/tasks/quadruped/quadruped.cc:33-221) with the
// ground ray cast it needs (Ground(), mjpc/utilities.cc:538-556 -> mj_ray against group-0 scenery).
//
// The ResidualFn member state the reference keeps between Transition() calls is a per-iteration
// snapshot in DevTask (filled by the host: mujoco_mpc_b200/task.py):
//   aux[0] mode   aux[1] mode_start_time   aux[2..3] position   aux[4..5] heading   aux[6] speed   aux[7] angvel
//   aux[8] gait   aux[9] phase_start   aux[10] phase_start_time   aux[11] phase_velocity   aux[12] ground
//   aux[13..16] orientation   aux[20..31] flip timing constants   aux[32..50] key "home"   aux[51..69] key "crouch"
//   iaux[0] torso body   iaux[1] head site   iaux[2] goal mocap   iaux[3..6] foot geoms FL,HL,FR,HR
//   iaux[7..11] parameter slots: biped type, heading, amplitude, duty ratio, flip direction
//   sensor_id[0..2] torso_subtreecom, torso_subtreelinvel, torso_angmom
#pragma once
#include "ps_step.cuh"

namespace ps {

// ---------------------------------------------------------------- ray casts against static scenery
PS_HD void ray_to_local(const double* pos, const double* mat, const double* pnt, const double* vec, double* lp, double* lv) {
  const double dif[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]};
  matT_vec3(lp, mat, dif); matT_vec3(lv, mat, vec);
}

PS_HD double ray_plane(const double* pos, const double* mat, const double* size, const double* pnt, const double* vec) {
  double lp[3], lv[3]; ray_to_local(pos, mat, pnt, vec, lp, lv);
  if (lv[2] > -kMinVal) return -1;
  const double x = -lp[2] / lv[2];
  if (x < 0) return -1;
  const double p0 = lp[0] + x * lv[0], p1 = lp[1] + x * lv[1];
  if ((size[0] <= 0 || fabs(p0) <= size[0]) && (size[1] <= 0 || fabs(p1) <= size[1])) return x;
  return -1;
}

PS_HD double ray_sphere(const double* pos, const double* size, const double* pnt, const double* vec) {
  const double dif[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]};
  const double a = dot3(vec, vec), b = dot3(vec, dif), c = dot3(dif, dif) - size[0] * size[0];
  double det = b * b - a * c;
  if (det < kMinVal) return -1;
  det = sqrt(det);
  const double x0 = (-b - det) / a, x1 = (-b + det) / a;
  if (x0 >= 0) return x0;
  if (x1 >= 0) return x1;
  return -1;
}

PS_HD double ray_box(const double* pos, const double* mat, const double* size, const double* pnt, const double* vec) {
  double lp[3], lv[3]; ray_to_local(pos, mat, pnt, vec, lp, lv);
  double x = -1;
  for (int i = 0; i < 3; i++) {
    if (fabs(lv[i]) <= kMinVal) continue;
    const int f0 = i == 0 ? 1 : 0, f1 = i == 2 ? 1 : 2;
    for (int side = -1; side <= 1; side += 2) {
      const double sol = (side * size[i] - lp[i]) / lv[i];
      if (sol < 0) continue;
      const double p0 = lp[f0] + sol * lv[f0], p1 = lp[f1] + sol * lv[f1];
      if (fabs(p0) <= size[f0] && fabs(p1) <= size[f1] && (x < 0 || sol < x)) x = sol;
    }
  }
  return x;
}

// height of the nearest group-0 geom under `pos`
template <class T, class D>
PS_NT double ground_height(const DevModel<T>& m, const D& d, const double* pos) {
  const double down[3] = {0, 0, -1}; const double offset = 0.5;
  const double query[3] = {pos[0], pos[1], pos[2] + offset};
  double dist = -1;
  for (int k = 0; k < m.nray; k++) {
    const int g = m.ray_geom[k];
    const double* gp = geom_world_pos(m, d, g); const double* gm = geom_world_mat(m, d, g); const double* sz = m.geom_size + 3 * g;
    const int tp = m.geom_type[g];
    const double x = tp == GEOM_PLANE ? ray_plane(gp, gm, sz, query, down)
                   : tp == GEOM_SPHERE ? ray_sphere(gp, sz, query, down) : ray_box(gp, gm, sz, query, down);
    if (x >= 0 && (x < dist || dist < 0)) dist = x;
  }
  return pos[2] + offset - dist;
}

// ---------------------------------------------------------------- helpers
namespace quad {
enum { MODE_QUADRUPED = 0, MODE_BIPED, MODE_WALK, MODE_SCRAMBLE, MODE_FLIP };
enum { FOOT_FL = 0, FOOT_HL, FOOT_FR, FOOT_HR };
enum { GAIT_TROT = 2 };
enum { A_MODE = 0, A_MODE_START, A_POSITION, A_HEADING = 4, A_SPEED = 6, A_ANGVEL, A_GAIT, A_PHASE_START, A_PHASE_START_TIME,
       A_PHASE_VELOCITY, A_GROUND, A_ORIENTATION, A_FLIP = 20, A_HOME = 32, A_CROUCH = 51 };
enum { F_CROUCH_TIME = 0, F_JUMP_TIME, F_FLIGHT_TIME, F_LAND_TIME, F_CROUCH_VEL, F_JUMP_ACC, F_JUMP_VEL, F_LAND_ACC, F_JUMP_ROT_VEL,
       F_JUMP_ROT_ACC, F_FLIGHT_ROT_VEL, F_LAND_ROT_ACC };
constexpr double kHeightQuadruped = 0.25, kHeightBiped = 0.6, kFootRadius = 0.02, kMinAngvel = 0.01, kLeapHeight = 0.5;

PS_HD double gait_phase(int gait, int foot) {
  // rows: stand, walk, trot, canter, gallop; columns FL HL FR HR (quadruped.h:77-85)
  const double tbl[20] = {0, 0, 0, 0, 0, 0.75, 0.5, 0.25, 0, 0.5, 0.5, 0, 0, 0.33, 0.33, 0.66, 0, 0.4, 0.05, 0.35};
  return tbl[4 * gait + foot];
}

PS_HD int as_int(double x) {
#ifdef __CUDA_ARCH__
  return (int)__double_as_longlong(x);
#else
  long long i; memcpy(&i, &x, sizeof(i)); return (int)i;
#endif
}

PS_HD void unit2(double* v) {
  const double n = sqrt(v[0] * v[0] + v[1] * v[1]);
  if (n < kMinVal) { v[0] = 1; v[1] = 0; } else { const double inv = 1 / n; v[0] *= inv; v[1] *= inv; }
}

PS_HD double step_height(double time, double footphase, double duty) {
  double angle = fmod(time + kPi - footphase, 2 * kPi) - kPi;
  double value = 0;
  if (duty < 1) { angle *= 0.5 / (1 - duty); value = cos_det(clampd(angle, -kPi / 2, kPi / 2)); }
  return fabs(value) < 1e-6 ? 0.0 : value;
}

PS_HD void walk_target(const double* a, double* pos, double time) {
  const double* position = a + A_POSITION; const double* heading = a + A_HEADING;
  if (fabs(a[A_ANGVEL]) < kMinAngvel) {
    double fwd[2] = {heading[0], heading[1]}; unit2(fwd);
    pos[0] = position[0] + heading[0] + time * a[A_SPEED] * fwd[0];
    pos[1] = position[1] + heading[1] + time * a[A_SPEED] * fwd[1];
  } else {
    double sn, cs; sincos_det(time * a[A_ANGVEL], &sn, &cs);
    pos[0] = cs * heading[0] + -sn * heading[1];
    pos[1] = sn * heading[0] + cs * heading[1];
    pos[0] += position[0]; pos[1] += position[1];
  }
}

PS_HD double flip_height(const double* a, double time) {
  const double* f = a + A_FLIP; const double ground = a[A_GROUND];
  if (time >= f[F_JUMP_TIME] + f[F_FLIGHT_TIME] + f[F_LAND_TIME]) return kHeightQuadruped + ground;
  double h = 0;
  if (time < f[F_JUMP_TIME]) h = kHeightQuadruped + time * f[F_CROUCH_VEL] + 0.5 * time * time * f[F_JUMP_ACC];
  else if (time >= f[F_JUMP_TIME] && time < f[F_JUMP_TIME] + f[F_FLIGHT_TIME]) {
    time -= f[F_JUMP_TIME]; h = kLeapHeight + f[F_JUMP_VEL] * time - 0.5 * 9.81 * time * time;
  } else if (time >= f[F_JUMP_TIME] + f[F_FLIGHT_TIME]) {
    time -= f[F_JUMP_TIME] + f[F_FLIGHT_TIME]; h = kLeapHeight - f[F_JUMP_VEL] * time + 0.5 * f[F_LAND_ACC] * time * time;
  }
  return h + ground;
}

PS_HD void flip_quat(const double* a, double* quat, double time, int flip_dir) {
  const double* f = a + A_FLIP;
  double angle = 0;
  if (time >= f[F_JUMP_TIME] + f[F_FLIGHT_TIME] + f[F_LAND_TIME]) angle = 2 * kPi;
  else if (time >= f[F_CROUCH_TIME] && time < f[F_JUMP_TIME]) {
    time -= f[F_CROUCH_TIME]; angle = 0.5 * f[F_JUMP_ROT_ACC] * time * time + f[F_JUMP_ROT_VEL] * time;
  } else if (time >= f[F_JUMP_TIME] && time < f[F_JUMP_TIME] + f[F_FLIGHT_TIME]) {
    time -= f[F_JUMP_TIME]; angle = kPi / 2 + f[F_FLIGHT_ROT_VEL] * time;
  } else if (time >= f[F_JUMP_TIME] + f[F_FLIGHT_TIME]) {
    time -= f[F_JUMP_TIME] + f[F_FLIGHT_TIME]; angle = 1.75 * kPi + f[F_FLIGHT_ROT_VEL] * time - 0.5 * f[F_LAND_ROT_ACC] * time * time;
  }
  const double axis[3] = {0, flip_dir ? 1.0 : -1.0, 0};
  double q[4]; axis_angle_quat(q, axis, angle);
  quat_mul(quat, a + A_ORIENTATION, q);
}

// mju_subQuat: rotation vector taking qb to qa
PS_HD void quat_difference(double* res, const double* qa, const double* qb) {
  const double qneg[4] = {qb[0], -qb[1], -qb[2], -qb[3]};
  double qd[4]; quat_mul(qd, qneg, qa);
  double axis[3] = {qd[1], qd[2], qd[3]};
  const double s = unit3(axis);
  double speed = 2 * atan2(s, qd[0]);
  if (speed > kPi) speed -= 2 * kPi;
  for (int k = 0; k < 3; k++) res[k] = axis[k] * speed;
}
}  // namespace quad

// ---------------------------------------------------------------- residual (42 values, 9 terms)
template <class T, class D>
PS_NT void quadruped_residual(const DevModel<T>& m, const D& d, const DevTask& t, double* r) {
  using namespace quad;
  const double* a = t.aux; const double* P = t.parameters;
  const int mode = (int)a[A_MODE], torso = t.iaux[0];
  const int handstand = as_int(P[t.iaux[7]]);
  const bool biped = mode == MODE_BIPED;
  const double* foot[4];
  for (int f = 0; f < 4; f++) foot[f] = geom_world_pos(m, d, t.iaux[3 + f]);
  double avg[3];
  if (biped) {
    const double* fa = handstand ? foot[FOOT_FL] : foot[FOOT_HL]; const double* fb = handstand ? foot[FOOT_FR] : foot[FOOT_HR];
    for (int k = 0; k < 3; k++) avg[k] = (fa[k] + fb[k]) * 0.5;
  } else {
    for (int k = 0; k < 3; k++) avg[k] = (foot[FOOT_HL][k] + foot[FOOT_HR][k] + foot[FOOT_FL][k] + foot[FOOT_FR][k]) * 0.25;
  }
  const double* xmat = d.xmat + 9 * torso;
  const double* goal = d.mocap_pos + 3 * t.iaux[2];
  const double* torso_pos = d.xipos + 3 * torso;
  const double mode_time = d.time - a[A_MODE_START];
  int c = 0;

  // Upright
  if (mode != MODE_FLIP) {
    r[c++] = biped ? xmat[6] - (handstand ? -1 : 1) : xmat[8] - 1;
    r[c++] = 0; r[c++] = 0;
  } else {
    double q[4]; flip_quat(a, q, mode_time, as_int(P[t.iaux[11]]));
    quat_difference(r + c, d.xquat + 4 * torso, q); c += 3;
  }

  // Height
  const double height_goal = biped ? kHeightBiped : kHeightQuadruped;
  if (mode == MODE_SCRAMBLE) r[c++] = 0;
  else if (mode == MODE_FLIP) r[c++] = torso_pos[2] - flip_height(a, mode_time);
  else r[c++] = (torso_pos[2] - avg[2]) - height_goal;

  // Position
  const double* head = d.site_xpos + 3 * t.iaux[1];
  double target[3] = {goal[0], goal[1], goal[2]};
  if (mode == MODE_WALK) walk_target(a, target, mode_time);
  r[c++] = head[0] - target[0];
  r[c++] = head[1] - target[1];
  r[c++] = mode == MODE_SCRAMBLE ? 2 * (head[2] - target[2]) : 0;

  // Gait
  const int gait = biped ? (int)GAIT_TROT : (int)a[A_GAIT];
  const double phase = a[A_PHASE_START] + (d.time - a[A_PHASE_START_TIME]) * a[A_PHASE_VELOCITY];
  const double amplitude = P[t.iaux[9]], duty = P[t.iaux[10]];
  for (int f = 0; f < 4; f++) {
    if (biped) {
      const bool front_hand = !handstand && (f == FOOT_FL || f == FOOT_FR);
      const bool back_hand = handstand && (f == FOOT_HL || f == FOOT_HR);
      if (front_hand || back_hand) { r[c++] = 0; continue; }
    }
    const double step = amplitude * step_height(phase, 2 * kPi * gait_phase(gait, f), duty);
    double query[3] = {foot[f][0], foot[f][1], foot[f][2]};
    if (mode == MODE_SCRAMBLE) {
      double v[3] = {goal[0] - foot[f][0], goal[1] - foot[f][1], 0};
      unit3(v);
      for (int k = 0; k < 3; k++) query[k] += v[k] * 0.15;
    }
    const double target_h = ground_height(m, d, query) + kFootRadius + step;
    double diff = foot[f][2] - target_h;
    if (mode == MODE_SCRAMBLE) diff = diff < 0 ? diff : 0;
    r[c++] = step ? diff : 0;
  }

  // Balance
  double com[3], vel[3];
  read_sensor(m, d, t.sensor_id[0], com); read_sensor(m, d, t.sensor_id[1], vel);
  const double fall_time = sqrt(2 * height_goal / 9.81);
  r[c++] = (com[0] + vel[0] * fall_time) - avg[0];
  r[c++] = (com[1] + vel[1] * fall_time) - avg[1];

  // Effort
  for (int i = 0; i < m.nu; i++) r[c + i] = d.actuator_force[i] * 2e-2;
  c += m.nu;

  // Posture
  const double* ref = a + A_HOME;
  bool free_legs = false;
  if (mode == MODE_FLIP) {
    const double* f = a + A_FLIP;
    if (mode_time < f[F_CROUCH_TIME]) ref = a + A_CROUCH;
    else if (mode_time < f[F_JUMP_TIME] + f[F_FLIGHT_TIME]) free_legs = true;
  }
  for (int i = 0; i < m.nu; i++) {
    double v = free_legs ? 0.0 : d.qpos[7 + i] - ref[7 + i];
    if (i % 3 == 0) v *= 2;                               // kJointPostureGain {2, 1, 1}
    r[c + i] = v;
  }
  if (biped) {
    if (handstand) { r[c + 4] *= 0.03; r[c + 5] *= 0.03; r[c + 10] *= 0.03; r[c + 11] *= 0.03; }
    else { r[c + 1] *= 0.03; r[c + 2] *= 0.03; r[c + 7] *= 0.03; r[c + 8] *= 0.03; }
  }
  c += m.nu;

  // Yaw
  double hd[2] = {xmat[0], xmat[3]};
  if (biped) { const int sg = handstand ? 1 : -1; hd[0] = sg * xmat[2]; hd[1] = sg * xmat[5]; }
  unit2(hd);
  double sn, cs; sincos_det(P[t.iaux[8]], &sn, &cs);
  r[c++] = hd[0] - cs; r[c++] = hd[1] - sn;

  // Angular momentum
  read_sensor(m, d, t.sensor_id[2], r + c);
}

}  // namespace ps
