// ORACLE (test infrastructure, not product).  Quadruped Flat residual, restating
// /root/reference/mjpc/tasks/quadruped/quadruped.cc:33-221 (Residual) with its helpers
// (:601-711 AverageFootPos / GetPhase / Walk / GetGait / StepHeight / FootStep / FlipHeight /
// FlipQuat) and Ground() (/root/reference/mjpc/utilities.cc:538-556), which calls MuJoCo's mj_ray.
// mj_ray (MuJoCo 3.1.4 engine_ray.c, not in the reference tree) is restated for the geom types the
// task's group-0 scenery uses: plane, sphere, box.
//
// The ResidualFn member state the reference keeps between Transition() calls arrives through the
// PsTask snapshot (include/ps_engine.h):
//   aux[0] current_mode_   aux[1] mode_start_time_   aux[2..3] position_   aux[4..5] heading_
//   aux[6] speed_          aux[7] angvel_            aux[8] gait index (current_gait_ as a number)
//   aux[9] phase_start_    aux[10] phase_start_time_ aux[11] phase_velocity_
//   aux[12] ground_        aux[13..16] orientation_
//   aux[20..31] flip constants: crouch_time, jump_time, flight_time, land_time, crouch_vel,
//               jump_acc, jump_vel, land_acc, jump_rot_vel, jump_rot_acc, flight_rot_vel, land_rot_acc
//   aux[32..50] key "home" qpos   aux[51..69] key "crouch" qpos
//   iaux[0] torso body  iaux[1] head site  iaux[2] goal mocap id  iaux[3..6] foot geoms FL,HL,FR,HR
//   iaux[7] biped-type parameter  iaux[8] Heading parameter  iaux[9] Amplitude  iaux[10] Duty ratio
//   iaux[11] Flip-dir parameter
//   sensor_id[0..2] torso_subtreecom, torso_subtreelinvel, torso_angmom
#include <cmath>
#include <cstdint>
#include <cstring>
#include "mm_math.h"
#include "mm_planner.h"

namespace mm {

// ---------------------------------------------------------------- mj_ray (plane / sphere / box)
static void rayMap(const double* pos, const double* mat, const double* pnt, const double* vec, double* lpnt, double* lvec) {
  const double dif[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]};
  mulMatTVec3(lpnt, mat, dif);
  mulMatTVec3(lvec, mat, vec);
}

static double rayPlane(const double* pos, const double* mat, const double* size, const double* pnt, const double* vec) {
  double lpnt[3], lvec[3]; rayMap(pos, mat, pnt, vec, lpnt, lvec);
  if (lvec[2] > -kMinVal) return -1;                 // not pointing at the front face
  const double x = -lpnt[2] / lvec[2];
  if (x < 0) return -1;
  const double p0 = lpnt[0] + x * lvec[0], p1 = lpnt[1] + x * lvec[1];
  if ((size[0] <= 0 || std::fabs(p0) <= size[0]) && (size[1] <= 0 || std::fabs(p1) <= size[1])) return x;
  return -1;
}

static double raySphere(const double* pos, const double* size, const double* pnt, const double* vec) {
  const double dif[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]};
  const double a = dot3(vec, vec), b = dot3(vec, dif), c = dot3(dif, dif) - size[0] * size[0];
  double det = b * b - a * c;                          // a x^2 + 2 b x + c = 0
  if (det < kMinVal) return -1;
  det = std::sqrt(det);
  const double x0 = (-b - det) / a, x1 = (-b + det) / a;
  if (x0 >= 0) return x0;
  if (x1 >= 0) return x1;
  return -1;
}

static double rayBox(const double* pos, const double* mat, const double* size, const double* pnt, const double* vec) {
  static const int face[3][2] = {{1, 2}, {0, 2}, {0, 1}};
  double lpnt[3], lvec[3]; rayMap(pos, mat, pnt, vec, lpnt, lvec);
  double x = -1;
  for (int i = 0; i < 3; i++) {
    if (std::fabs(lvec[i]) <= kMinVal) continue;
    for (int side = -1; side <= 1; side += 2) {
      const double sol = (side * size[i] - lpnt[i]) / lvec[i];
      if (sol < 0) continue;
      const double p0 = lpnt[face[i][0]] + sol * lvec[face[i][0]], p1 = lpnt[face[i][1]] + sol * lvec[face[i][1]];
      if (std::fabs(p0) <= size[face[i][0]] && std::fabs(p1) <= size[face[i][1]] && (x < 0 || sol < x)) x = sol;
    }
  }
  return x;
}

double ray(const Model* m, const Data* d, const double pnt[3], const double vec[3], const unsigned char* geomgroup,
           int flg_static, int bodyexclude, int* geomid) {
  double dist = -1; int id = -1;
  for (int g = 0; g < m->ngeom; g++) {
    const int body = m->geom_bodyid[g];
    if (body == bodyexclude) continue;
    int weld = body; while (weld > 0 && m->body_dofnum[weld] == 0) weld = m->body_parentid[weld];
    if (!flg_static && weld == 0) continue;
    if (geomgroup) { int grp = m->geom_group[g]; grp = grp < 0 ? 0 : (grp > 5 ? 5 : grp); if (!geomgroup[grp]) continue; }
    const double* pos = &d->geom_xpos[3 * g]; const double* mat = &d->geom_xmat[9 * g]; const double* size = m->geom_size + 3 * g;
    double x = -1;
    switch (m->geom_type[g]) {
      case PS_GEOM_PLANE: x = rayPlane(pos, mat, size, pnt, vec); break;
      case PS_GEOM_SPHERE: x = raySphere(pos, size, pnt, vec); break;
      case PS_GEOM_BOX: x = rayBox(pos, mat, size, pnt, vec); break;
      default: break;                                  // other types never sit in group 0 for this task
    }
    if (x >= 0 && (x < dist || dist < 0)) { dist = x; id = g; }
  }
  if (geomid) *geomid = id;
  return dist;
}

// utilities.cc:538-556
static double ground(const Model* m, const Data* d, const double pos[3]) {
  const double down[3] = {0, 0, -1}; const double offset = 0.5;
  const unsigned char group[6] = {1, 0, 0, 0, 0, 0};
  const double query[3] = {pos[0], pos[1], pos[2] + offset};
  const double dist = ray(m, d, query, down, group, 1, -1, nullptr);
  return pos[2] + offset - dist;
}

// ---------------------------------------------------------------- residual helpers
namespace {
enum { kModeQuadruped = 0, kModeBiped, kModeWalk, kModeScramble, kModeFlip };
enum { kFootFL = 0, kFootHL, kFootFR, kFootHR, kNumFoot };
enum { kGaitStand = 0, kGaitWalk, kGaitTrot, kGaitCanter, kGaitGallop };
const double kGaitPhase[5][4] = {{0, 0, 0, 0}, {0, 0.75, 0.5, 0.25}, {0, 0.5, 0.5, 0}, {0, 0.33, 0.33, 0.66}, {0, 0.4, 0.05, 0.35}};
const double kHeightQuadruped = 0.25, kHeightBiped = 0.6, kFootRadius = 0.02, kMinAngvel = 0.01, kLeapHeight = 0.5;
const double kJointPostureGain[3] = {2, 1, 1};

int asInt(double x) { int64_t i; std::memcpy(&i, &x, sizeof(i)); return (int)i; }   // utilities.cc ReinterpretAsInt

struct Snapshot {
  int mode; double mode_start_time, position[2], heading[2], speed, angvel; int gait;
  double phase_start, phase_start_time, phase_velocity, ground, orientation[4];
  double crouch_time, jump_time, flight_time, land_time, crouch_vel, jump_acc, jump_vel, land_acc, jump_rot_vel, jump_rot_acc,
      flight_rot_vel, land_rot_acc;
  const double* home; const double* crouch;
};

Snapshot unpack(const PsTask* t) {
  const double* a = t->aux; Snapshot s;
  s.mode = (int)a[0]; s.mode_start_time = a[1]; s.position[0] = a[2]; s.position[1] = a[3]; s.heading[0] = a[4]; s.heading[1] = a[5];
  s.speed = a[6]; s.angvel = a[7]; s.gait = (int)a[8]; s.phase_start = a[9]; s.phase_start_time = a[10]; s.phase_velocity = a[11];
  s.ground = a[12]; for (int k = 0; k < 4; k++) s.orientation[k] = a[13 + k];
  const double* f = a + 20;
  s.crouch_time = f[0]; s.jump_time = f[1]; s.flight_time = f[2]; s.land_time = f[3]; s.crouch_vel = f[4]; s.jump_acc = f[5];
  s.jump_vel = f[6]; s.land_acc = f[7]; s.jump_rot_vel = f[8]; s.jump_rot_acc = f[9]; s.flight_rot_vel = f[10]; s.land_rot_acc = f[11];
  s.home = a + 32; s.crouch = a + 51;
  return s;
}

// quadruped.cc:633-648
void walkTarget(const Snapshot& s, double pos[2], double time) {
  if (std::fabs(s.angvel) < kMinAngvel) {
    double fwd[2] = {s.heading[0], s.heading[1]};
    const double n = std::sqrt(fwd[0] * fwd[0] + fwd[1] * fwd[1]);
    if (n < kMinVal) { fwd[0] = 1; fwd[1] = 0; } else { const double inv = 1 / n; fwd[0] *= inv; fwd[1] *= inv; }
    pos[0] = s.position[0] + s.heading[0] + time * s.speed * fwd[0];
    pos[1] = s.position[1] + s.heading[1] + time * s.speed * fwd[1];
  } else {
    const double angle = time * s.angvel, c = cosDet(angle), sn = sinDet(angle);
    pos[0] = c * s.heading[0] + -sn * s.heading[1];
    pos[1] = sn * s.heading[0] + c * s.heading[1];
    pos[0] += s.position[0]; pos[1] += s.position[1];
  }
}

// quadruped.cc:658-667
double stepHeight(double time, double footphase, double duty_ratio) {
  double angle = std::fmod(time + kPi - footphase, 2 * kPi) - kPi;
  double value = 0;
  if (duty_ratio < 1) {
    angle *= 0.5 / (1 - duty_ratio);
    value = cosDet(clip(angle, -kPi / 2, kPi / 2));
  }
  return std::fabs(value) < 1e-6 ? double(0.0) : value;
}

// quadruped.cc:681-697
double flipHeight(const Snapshot& s, double time) {
  if (time >= s.jump_time + s.flight_time + s.land_time) return kHeightQuadruped + s.ground;
  double h = 0;
  if (time < s.jump_time) h = kHeightQuadruped + time * s.crouch_vel + 0.5 * time * time * s.jump_acc;
  else if (time >= s.jump_time && time < s.jump_time + s.flight_time) { time -= s.jump_time; h = kLeapHeight + s.jump_vel * time - 0.5 * 9.81 * time * time; }
  else if (time >= s.jump_time + s.flight_time) { time -= s.jump_time + s.flight_time; h = kLeapHeight - s.jump_vel * time + 0.5 * s.land_acc * time * time; }
  return h + s.ground;
}

// quadruped.cc:702-722
void flipQuat(const Snapshot& s, double quat[4], double time, int flip_dir) {
  double angle = 0;
  if (time >= s.jump_time + s.flight_time + s.land_time) angle = 2 * kPi;
  else if (time >= s.crouch_time && time < s.jump_time) { time -= s.crouch_time; angle = 0.5 * s.jump_rot_acc * time * time + s.jump_rot_vel * time; }
  else if (time >= s.jump_time && time < s.jump_time + s.flight_time) { time -= s.jump_time; angle = kPi / 2 + s.flight_rot_vel * time; }
  else if (time >= s.jump_time + s.flight_time) { time -= s.jump_time + s.flight_time; angle = 1.75 * kPi + s.flight_rot_vel * time - 0.5 * s.land_rot_acc * time * time; }
  const double axis[3] = {0, flip_dir ? 1.0 : -1.0, 0};
  double q[4]; axisAngle2Quat(q, axis, angle);
  mulQuat(quat, s.orientation, q);
}

// mju_subQuat: 3D velocity taking qb to qa
void subQuat(double res[3], const double qa[4], const double qb[4]) {
  const double qneg[4] = {qb[0], -qb[1], -qb[2], -qb[3]};
  double qdif[4]; mulQuat(qdif, qneg, qa);
  double axis[3] = {qdif[1], qdif[2], qdif[3]};
  const double sin_a_2 = normalize3(axis);
  double speed = 2 * std::atan2(sin_a_2, qdif[0]);
  if (speed > kPi) speed -= 2 * kPi;
  for (int k = 0; k < 3; k++) res[k] = axis[k] * speed;
}
}  // namespace

void quadrupedResidual(const Model* m, const Data* d, const PsTask* t, double* residual) {
  const Snapshot s = unpack(t);
  const int torso = t->iaux[0], head_site = t->iaux[1], goal_mocap = t->iaux[2];
  const double* P = t->parameters;
  const int biped_handstand = asInt(P[t->iaux[7]]);
  int counter = 0;
  const double* foot_pos[kNumFoot];
  for (int f = 0; f < kNumFoot; f++) foot_pos[f] = &d->geom_xpos[3 * t->iaux[3 + f]];

  // average foot position (quadruped.cc:601-618)
  double avg[3];
  if (s.mode == kModeBiped) {
    const double* a = biped_handstand ? foot_pos[kFootFL] : foot_pos[kFootHL];
    const double* b = biped_handstand ? foot_pos[kFootFR] : foot_pos[kFootHR];
    for (int k = 0; k < 3; k++) avg[k] = (a[k] + b[k]) * 0.5;
  } else {
    for (int k = 0; k < 3; k++) avg[k] = (foot_pos[kFootHL][k] + foot_pos[kFootHR][k] + foot_pos[kFootFL][k] + foot_pos[kFootFR][k]) * 0.25;
  }
  const double* torso_xmat = &d->xmat[9 * torso];
  const double* goal_pos = &d->mocap_pos[3 * goal_mocap];
  const double* compos = &d->sensordata[m->sensor_adr[t->sensor_id[0]]];

  // ---------- Upright ----------
  if (s.mode != kModeFlip) {
    if (s.mode == kModeBiped) residual[counter++] = torso_xmat[6] - (biped_handstand ? -1 : 1);
    else residual[counter++] = torso_xmat[8] - 1;
    residual[counter++] = 0; residual[counter++] = 0;
  } else {
    double quat[4]; flipQuat(s, quat, d->time - s.mode_start_time, asInt(P[t->iaux[11]]));
    subQuat(residual + counter, &d->xquat[4 * torso], quat);
    counter += 3;
  }

  // ---------- Height ----------
  const double* torso_pos = &d->xipos[3 * torso];
  const bool is_biped = s.mode == kModeBiped;
  const double height_goal = is_biped ? kHeightBiped : kHeightQuadruped;
  if (s.mode == kModeScramble) residual[counter++] = 0;
  else if (s.mode == kModeFlip) residual[counter++] = torso_pos[2] - flipHeight(s, d->time - s.mode_start_time);
  else residual[counter++] = (torso_pos[2] - avg[2]) - height_goal;

  // ---------- Position ----------
  const double* head = &d->site_xpos[3 * head_site];
  double target[3] = {0, 0, 0};
  if (s.mode == kModeWalk) walkTarget(s, target, d->time - s.mode_start_time);
  else { target[0] = goal_pos[0]; target[1] = goal_pos[1]; target[2] = goal_pos[2]; }
  residual[counter++] = head[0] - target[0];
  residual[counter++] = head[1] - target[1];
  residual[counter++] = s.mode == kModeScramble ? 2 * (head[2] - target[2]) : double(0);

  // ---------- Gait ----------
  const int gait = s.mode == kModeBiped ? kGaitTrot : s.gait;
  double step[kNumFoot];
  {
    const double phase = s.phase_start + (d->time - s.phase_start_time) * s.phase_velocity;
    const double amplitude = P[t->iaux[9]], duty = P[t->iaux[10]];
    for (int f = 0; f < kNumFoot; f++) step[f] = amplitude * stepHeight(phase, 2 * kPi * kGaitPhase[gait][f], duty);
  }
  for (int f = 0; f < kNumFoot; f++) {
    if (is_biped) {
      const bool front_hand = !biped_handstand && (f == kFootFL || f == kFootFR);
      const bool back_hand = biped_handstand && (f == kFootHL || f == kFootHR);
      if (front_hand || back_hand) { residual[counter++] = 0; continue; }
    }
    double query[3] = {foot_pos[f][0], foot_pos[f][1], foot_pos[f][2]};
    if (s.mode == kModeScramble) {
      double v[3];
      for (int k = 0; k < 3; k++) v[k] = goal_pos[k] - torso_pos[k];
      normalize3(v);
      for (int k = 0; k < 3; k++) v[k] = goal_pos[k] - foot_pos[f][k];
      v[2] = 0; normalize3(v);
      for (int k = 0; k < 3; k++) query[k] += v[k] * 0.15;
    }
    const double ground_height = ground(m, d, query);
    const double height_target = ground_height + kFootRadius + step[f];
    double diff = foot_pos[f][2] - height_target;
    if (s.mode == kModeScramble) diff = diff < 0 ? diff : double(0);      // mju_min(0, diff)
    residual[counter++] = step[f] ? diff : double(0);
  }

  // ---------- Balance ----------
  const double* comvel = &d->sensordata[m->sensor_adr[t->sensor_id[1]]];
  const double fall_time = std::sqrt(2 * height_goal / 9.81);
  residual[counter++] = (compos[0] + comvel[0] * fall_time) - avg[0];
  residual[counter++] = (compos[1] + comvel[1] * fall_time) - avg[1];

  // ---------- Effort ----------
  for (int i = 0; i < m->nu; i++) residual[counter + i] = d->actuator_force[i] * 2e-2;
  counter += m->nu;

  // ---------- Posture ----------
  for (int i = 0; i < m->nu; i++) residual[counter + i] = d->qpos[7 + i] - s.home[7 + i];
  if (s.mode == kModeFlip) {
    const double flip_time = d->time - s.mode_start_time;
    if (flip_time < s.crouch_time) for (int i = 0; i < m->nu; i++) residual[counter + i] = d->qpos[7 + i] - s.crouch[7 + i];
    else if (flip_time >= s.crouch_time && flip_time < s.jump_time + s.flight_time) for (int i = 0; i < m->nu; i++) residual[counter + i] = 0;
  }
  for (int f = 0; f < kNumFoot; f++) for (int j = 0; j < 3; j++) residual[counter + 3 * f + j] *= kJointPostureGain[j];
  if (s.mode == kModeBiped) {
    static const int loose_hand[4] = {4, 5, 10, 11}, loose_foot[4] = {1, 2, 7, 8};
    for (int k = 0; k < 4; k++) residual[counter + (biped_handstand ? loose_hand[k] : loose_foot[k])] *= 0.03;
  }
  counter += m->nu;

  // ---------- Yaw ----------
  double hd[2] = {torso_xmat[0], torso_xmat[3]};
  if (s.mode == kModeBiped) { const int sg = biped_handstand ? 1 : -1; hd[0] = sg * torso_xmat[2]; hd[1] = sg * torso_xmat[5]; }
  {
    const double n = std::sqrt(hd[0] * hd[0] + hd[1] * hd[1]);       // mju_normalize(., 2)
    if (n < kMinVal) { hd[0] = 1; hd[1] = 0; } else { const double inv = 1 / n; hd[0] *= inv; hd[1] *= inv; }
  }
  const double heading_goal = P[t->iaux[8]];
  residual[counter++] = hd[0] - cosDet(heading_goal);
  residual[counter++] = hd[1] - sinDet(heading_goal);

  // ---------- Angular momentum ----------
  const double* angmom = &d->sensordata[m->sensor_adr[t->sensor_id[2]]];
  for (int k = 0; k < 3; k++) residual[counter++] = angmom[k];
}

}  // namespace mm
