// ORACLE (test infrastructure, not product).  CPU FP64 restatement of the MuJoCo 3.1.4
// forward step (mj_step / mj_forward) for the model class the six benchmarked MJPC tasks use.
// The reference calls into MuJoCo at /root/reference/mjpc/trajectory.cc:158 (mj_step) and :198
// (mj_forward); MuJoCo itself is a third-party dependency pinned at
// /root/reference/CMakeLists.txt:58-61 (tag 3.1.4) and is NOT present in /root/reference, so its
// published algorithm is restated here stage by stage (fwdPosition -> sensorPos -> fwdVelocity
// -> sensorVel -> fwdActuation -> fwdAcceleration -> fwdConstraint -> sensorAcc -> integrate).
// Parity against real MuJoCo: UNPINNED (no MuJoCo binary or golden mj_step vectors exist in the
// reference tree).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may use it.
#ifndef ORACLE_MM_ENGINE_H_
#define ORACLE_MM_ENGINE_H_
#include <vector>
#include "../include/ps_model.h"

namespace mm {

typedef PsModel Model;

enum { CNSTR_FRICTION_DOF = 0, CNSTR_LIMIT_JOINT, CNSTR_CONTACT_FRICTIONLESS,
       CNSTR_CONTACT_PYRAMIDAL, CNSTR_CONTACT_ELLIPTIC };
enum { STATE_SATISFIED = 0, STATE_QUADRATIC, STATE_LINEARNEG, STATE_LINEARPOS, STATE_CONE };
enum { WARN_BADQPOS = 0, WARN_BADQVEL, WARN_BADQACC, WARN_BADCTRL, N_WARN };

struct Contact {
  double dist, pos[3], frame[9], includemargin, friction[5], solref[2], solimp[5], mu;
  double H[36];
  int dim, geom1, geom2, efc_address;
};

struct Data {
  int nefc = 0, ncon = 0;
  double time = 0;
  int warning[N_WARN] = {0, 0, 0, 0};
  int solver_iter = 0;
  std::vector<double> qpos, qvel, act, ctrl, qacc, qacc_warmstart, act_dot, mocap_pos, mocap_quat,
      userdata, sensordata, xfrc_applied;   // xfrc_applied: [force; torque] per body, world frame (mjData.xfrc_applied)
  // position dependent
  std::vector<double> xpos, xquat, xmat, xipos, ximat, xanchor, xaxis, geom_xpos, geom_xmat,
      site_xpos, site_xmat, subtree_com, cinert, crb, cdof, qM, qLD, qLDiagInv,
      actuator_length;
  // velocity dependent
  std::vector<double> cvel, cdof_dot, qfrc_bias, qfrc_spring, qfrc_damper, qfrc_fluid,
      qfrc_passive, actuator_velocity, subtree_linvel, subtree_angmom;
  // force / acceleration
  std::vector<double> actuator_force, qfrc_actuator, qfrc_smooth, qacc_smooth, qfrc_constraint,
      cacc, cfrc_body, qDeriv;
  // constraints
  std::vector<Contact> contact;
  std::vector<double> efc_J, efc_pos, efc_margin, efc_frictionloss, efc_diagApprox, efc_KBIP,
      efc_D, efc_R, efc_vel, efc_aref, efc_b, efc_force;
  std::vector<int> efc_type, efc_id, efc_state;
};

// residual callback, evaluated at the end of the acceleration stage (what mjcb_sensor does in
// the reference: /root/reference/mjpc/testspeed.cc:36-40).  Writes sensordata[0:nresidual].
typedef void (*SensorCallback)(const Model* m, Data* d, void* user);

void makeData(const Model* m, Data* d);
void resetData(const Model* m, Data* d);
void forward(const Model* m, Data* d, SensorCallback cb = nullptr, void* user = nullptr);
void step(const Model* m, Data* d, SensorCallback cb = nullptr, void* user = nullptr);

// stages (exposed for unit tests)
void fwdPosition(const Model* m, Data* d);
void fwdVelocity(const Model* m, Data* d);
void fwdActuation(const Model* m, Data* d);
void fwdAcceleration(const Model* m, Data* d);
void fwdConstraint(const Model* m, Data* d);
void sensorPos(const Model* m, Data* d);
void sensorVel(const Model* m, Data* d);
void sensorAcc(const Model* m, Data* d);
void rne(const Model* m, Data* d, const double* qvel_or_null, double* result);
void mulM(const Model* m, const Data* d, double* res, const double* vec);
void solveM(const Model* m, const Data* d, double* x);
void jac(const Model* m, const Data* d, double* jacp, double* jacr, const double point[3], int body);
double ray(const Model* m, const Data* d, const double pnt[3], const double vec[3],
           const unsigned char* geomgroup, int flg_static, int bodyexclude, int* geomid);
int name2id(const Model* m, int objtype, const char* name);

extern long long g_dbg_line_evals;
}  // namespace mm
#endif
