// ORACLE BUILD SHIM (test infrastructure).  Stand-in for <mujoco/mujoco.h> (MuJoCo 3.1.4, a third-party dependency the
// reference fetches at configure time and that is absent here) so that the reference's OWN sources -- mjpc/norm.cc,
// spline/spline.cc, task.cc, trajectory.cc, utilities.cc, states/state.cc, planners/planner.cc,
// planners/sampling/{planner,policy}.cc and the six tasks/*/*.cc -- compile UNMODIFIED from /root/reference into
// oracle/_ref/ (recipe: oracle/Makefile).  It declares the subset of the public API those files use, with MuJoCo's
// names, types and enum values.  mjModel / mjData are VIEWS: mjModel is filled from a PsModel (include/ps_model.h, which
// mirrors mjModel field for field), mjData's arrays alias the oracle's mm::Data, and mj_step / mj_forward / mj_ray /
// mj_name2id run the oracle's restated MuJoCo step (oracle/ref_glue.cc).  So everything ABOVE mj_step is the
// reference's own code; the physics below it stays the restatement (parity against real MuJoCo: UNPINNED).
// The mju_* helpers are inline restatements of engine_util_{blas,spatial,misc} with MuJoCo's operation order.
#ifndef ORACLE_SHIM_MUJOCO_H_
#define ORACLE_SHIM_MUJOCO_H_
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

typedef double mjtNum;
typedef unsigned char mjtByte;
#define mjMINVAL 1E-15
#define mjMAXVAL 1E+10
#define mjPI 3.14159265358979323846
#define mjMINMU 1E-5
#define mjMAXLINE 100
#define mjMAXLINEPNT 1000
#define mjMAXPLANEGRID 200
#define mjNGROUP 6
#define mjMAXUISECT 10
#define mjMAXUIITEM 200
#define mjMAXUITEXT 300
#define mjMAXUINAME 40
#define mjMAXUIMULTI 35
#define mjMAXUIEDIT 7
#define mjMAXUIRECT 25
#define mjNEQDATA 11
#define mjNDYN 10
#define mjNGAIN 10
#define mjNBIAS 10
#define mjMAX(a, b) (((a) > (b)) ? (a) : (b))
#define mjMIN(a, b) (((a) < (b)) ? (a) : (b))
#define mjVERSION_HEADER 314

// ---- enums (values of MuJoCo 3.1.4's mjmodel.h / mjdata.h / mjvisualize.h / mjui.h)
typedef enum { mjOBJ_UNKNOWN = 0, mjOBJ_BODY, mjOBJ_XBODY, mjOBJ_JOINT, mjOBJ_DOF, mjOBJ_GEOM, mjOBJ_SITE, mjOBJ_CAMERA, mjOBJ_LIGHT,
               mjOBJ_FLEX, mjOBJ_MESH, mjOBJ_SKIN, mjOBJ_HFIELD, mjOBJ_TEXTURE, mjOBJ_MATERIAL, mjOBJ_PAIR, mjOBJ_EXCLUDE,
               mjOBJ_EQUALITY, mjOBJ_TENDON, mjOBJ_ACTUATOR, mjOBJ_SENSOR, mjOBJ_NUMERIC, mjOBJ_TEXT, mjOBJ_TUPLE, mjOBJ_KEY,
               mjOBJ_PLUGIN } mjtObj;
typedef enum { mjGEOM_PLANE = 0, mjGEOM_HFIELD, mjGEOM_SPHERE, mjGEOM_CAPSULE, mjGEOM_ELLIPSOID, mjGEOM_CYLINDER, mjGEOM_BOX,
               mjGEOM_MESH, mjGEOM_SDF, mjNGEOMTYPES, mjGEOM_ARROW = 100, mjGEOM_ARROW1, mjGEOM_ARROW2, mjGEOM_LINE,
               mjGEOM_LINEBOX, mjGEOM_FLEX, mjGEOM_SKIN, mjGEOM_LABEL, mjGEOM_TRIANGLE, mjGEOM_NONE = 1001 } mjtGeom;
typedef enum { mjJNT_FREE = 0, mjJNT_BALL, mjJNT_SLIDE, mjJNT_HINGE } mjtJoint;
typedef enum { mjINT_EULER = 0, mjINT_RK4, mjINT_IMPLICIT, mjINT_IMPLICITFAST } mjtIntegrator;
typedef enum { mjSTAGE_NONE = 0, mjSTAGE_POS, mjSTAGE_VEL, mjSTAGE_ACC } mjtStage;
// sensor types: the numbering of include/ps_model.h (PS_SENS_*), which is what the flat models carry in sensor_type
typedef enum { mjSENS_TOUCH = 0, mjSENS_ACCELEROMETER = 1, mjSENS_JOINTPOS = 8, mjSENS_JOINTVEL = 9, mjSENS_FRAMEPOS = 25, mjSENS_FRAMEQUAT = 26,
               mjSENS_FRAMEXAXIS = 27, mjSENS_FRAMEYAXIS = 28, mjSENS_FRAMEZAXIS = 29, mjSENS_FRAMELINVEL = 30, mjSENS_FRAMEANGVEL = 31,
               mjSENS_FRAMELINACC = 32, mjSENS_FRAMEANGACC = 33, mjSENS_SUBTREECOM = 34, mjSENS_SUBTREELINVEL = 35,
               mjSENS_SUBTREEANGMOM = 36, mjSENS_USER = 39 } mjtSensor;
typedef enum { mjWARN_INERTIA = 0, mjWARN_CONTACTFULL, mjWARN_CNSTRFULL, mjWARN_VGEOMFULL, mjWARN_BADQPOS, mjWARN_BADQVEL,
               mjWARN_BADQACC, mjWARN_BADCTRL, mjNWARNING } mjtWarning;
typedef enum { mjCAT_STATIC = 1, mjCAT_DYNAMIC = 2, mjCAT_DECOR = 4, mjCAT_ALL = 7 } mjtCatBit;
typedef enum { mjITEM_END = -2, mjITEM_SECTION = -1, mjITEM_SEPARATOR = 0, mjITEM_STATIC, mjITEM_BUTTON, mjITEM_CHECKINT,
               mjITEM_CHECKBYTE, mjITEM_RADIO, mjITEM_RADIOLINE, mjITEM_SELECT, mjITEM_SLIDERINT, mjITEM_SLIDERNUM,
               mjITEM_EDITINT, mjITEM_EDITNUM, mjITEM_EDITFLOAT, mjITEM_EDITTXT, mjNITEM } mjtItem;

// ---- model / data views
struct mjOption { mjtNum timestep, impratio, tolerance, ls_tolerance, density, viscosity; mjtNum gravity[3], wind[3];
                  int integrator, cone, solver, iterations, ls_iterations, disableflags, enableflags; };
struct mjStatistic { mjtNum meaninertia; };
struct mjWarningStat { int lastinfo, number; };
struct PsModel_;
struct mjModel_ {
  mjOption opt; mjStatistic stat;
#include "mjmodel_fields.inc"          // every array / size PsModel mirrors, under mjModel's names and types
  mjtNum *key_mpos, *key_mquat;
  const PsModel_* ps;          // shim: the flat model this view was filled from (what the oracle physics reads)
};
typedef struct mjModel_ mjModel;
struct mjData_ {
  mjtNum time;
  int nefc, ncon;
  mjWarningStat warning[mjNWARNING];
  mjtNum *qpos, *qvel, *act, *ctrl, *qacc, *qacc_warmstart, *mocap_pos, *mocap_quat, *userdata, *sensordata, *xfrc_applied;
  mjtNum *xpos, *xquat, *xmat, *xipos, *ximat, *geom_xpos, *geom_xmat, *site_xpos, *site_xmat, *subtree_com, *actuator_force;
  void* shim;                  // shim: the mm::Data behind the views
};
typedef struct mjData_ mjData;

// ---- visualisation / UI types: declared so that drawing helpers compile; nothing on the planning path runs them
struct mjvGeom_ { int type, dataid, objtype, objid, category, texid, texuniform, texcoord, segid; float texrepeat[2], size[3], pos[3],
                  mat[9], rgba[4], emission, specular, shininess, reflectance; char label[100]; float camdist, modelrbound;
                  mjtByte transparent; };
typedef struct mjvGeom_ mjvGeom;
struct mjvScene_ { int maxgeom, ngeom; mjvGeom* geoms; };
typedef struct mjvScene_ mjvScene;
struct mjvFigure_ {
  int flg_legend, flg_ticklabel[2], flg_extend, flg_barplot, flg_selection, flg_symmetric;
  float linewidth, gridwidth; int gridsize[2]; float gridrgb[3], figurergba[4], panergba[4], legendrgba[4], textrgb[3],
      linergb[mjMAXLINE][3], range[2][2];
  char xformat[20], yformat[20], minwidth[20], title[1000], xlabel[100], linename[mjMAXLINE][100];
  int legendoffset, subplot, highlight[2], highlightid; float selection; int linepnt[mjMAXLINE];
  float linedata[mjMAXLINE][2 * mjMAXLINEPNT]; int xaxispixel[2], yaxispixel[2]; float xaxisdata[2], yaxisdata[2];
};
typedef struct mjvFigure_ mjvFigure;
struct mjrRect_ { int left, bottom, width, height; };
typedef struct mjrRect_ mjrRect;
struct mjuiDef_ { int type; char name[mjMAXUINAME]; int state; void* pdata; char other[mjMAXUITEXT]; };
typedef struct mjuiDef_ mjuiDef;
struct mjUI_ { int nsect; };
typedef struct mjUI_ mjUI;
struct mjuiState_ { int nrect; };
typedef struct mjuiState_ mjuiState;
struct mjrContext_ { int unused; };
typedef struct mjrContext_ mjrContext;
struct mjvOption_ { int unused; };
typedef struct mjvOption_ mjvOption;
struct mjvCamera_ { int unused; };
typedef struct mjvCamera_ mjvCamera;
struct mjvPerturb_ { int unused; };
typedef struct mjvPerturb_ mjvPerturb;

typedef void (*mjfSensor)(const mjModel* m, mjData* d, int stage);
typedef void (*mjfGeneric)(const mjModel* m, mjData* d);
extern mjfSensor mjcb_sensor;
extern mjfGeneric mjcb_control;

// ---- engine entry points (oracle/ref_glue.cc)
mjData* mj_makeData(const mjModel* m);
void mj_deleteData(mjData* d);
void mj_deleteModel(mjModel* m);
mjModel* mj_copyModel(mjModel* dest, const mjModel* src);
mjData* mj_copyData(mjData* dest, const mjModel* m, const mjData* src);
void mj_resetData(const mjModel* m, mjData* d);
void mj_resetDataKeyframe(const mjModel* m, mjData* d, int key);
void mj_step(const mjModel* m, mjData* d);
void mj_forward(const mjModel* m, mjData* d);
void mj_kinematics(const mjModel* m, mjData* d);
int mj_name2id(const mjModel* m, int type, const char* name);
const char* mj_id2name(const mjModel* m, int type, int id);
mjtNum mj_ray(const mjModel* m, const mjData* d, const mjtNum pnt[3], const mjtNum vec[3], const mjtByte* geomgroup,
              mjtByte flg_static, int bodyexclude, int geomid[1]);
void mj_differentiatePos(const mjModel* m, mjtNum* qvel, mjtNum dt, const mjtNum* qpos1, const mjtNum* qpos2);
void mj_integratePos(const mjModel* m, mjtNum* qpos, const mjtNum* qvel, mjtNum dt);
void mj_jac(const mjModel* m, const mjData* d, mjtNum* jacp, mjtNum* jacr, const mjtNum point[3], int body);
mjModel* mj_loadXML(const char* filename, const void* vfs, char* error, int error_sz);
void mjv_initGeom(mjvGeom* geom, int type, const mjtNum size[3], const mjtNum pos[3], const mjtNum mat[9], const float rgba[4]);
void mjv_makeConnector(mjvGeom* geom, int type, mjtNum width, mjtNum a0, mjtNum a1, mjtNum a2, mjtNum b0, mjtNum b1, mjtNum b2);
void mjui_add(mjUI* ui, const mjuiDef* def);
void mjr_figure(mjrRect viewport, mjvFigure* fig, const mjrContext* con);

// ---- mju_* (engine_util_*), inline; summation / operation order as in MuJoCo
// MuJoCo routes mju_error through a user handler when one is installed (mju_user_error) and terminates the process
// otherwise.  Same here: no handler (the default -- every test and every golden vector) aborts; the timed reference arm of
// bench.py installs a counting handler (ref_planner_wrap.cc: refp_tolerate_errors) so that one rollout the reference has no
// answer for -- Ground() with a foot more than 0.5 m under the floor, utilities.cc:551 "SHOULD NOT OCCUR" -- does not end the run.
typedef void (*mjfShimErrorHook)(const char*);
inline mjfShimErrorHook& mju_shim_error_hook() { static mjfShimErrorHook h = nullptr; return h; }
inline void mju_error(const char* msg, ...) {
  char buf[512]; va_list a; va_start(a, msg); std::vsnprintf(buf, sizeof(buf), msg, a); va_end(a);
  if (mjfShimErrorHook h = mju_shim_error_hook()) { h(buf); return; }
  std::fprintf(stderr, "mju_error: %s\n", buf); std::abort();
}
inline void mju_error_i(const char* msg, int i) { std::fprintf(stderr, "mju_error: "); std::fprintf(stderr, msg, i); std::fprintf(stderr, "\n"); std::abort(); }
inline void mju_error_s(const char* msg, const char* s) { std::fprintf(stderr, "mju_error: "); std::fprintf(stderr, msg, s); std::fprintf(stderr, "\n"); std::abort(); }
inline void mju_warning(const char* msg, ...) { va_list a; va_start(a, msg); std::fprintf(stderr, "mju_warning: "); std::vfprintf(stderr, msg, a); std::fprintf(stderr, "\n"); va_end(a); }
inline void mju_warning_i(const char* msg, int i) { std::fprintf(stderr, msg, i); std::fprintf(stderr, "\n"); }
inline void mju_warning_s(const char* msg, const char* s) { std::fprintf(stderr, msg, s); std::fprintf(stderr, "\n"); }
inline void* mju_malloc(size_t n) { return std::malloc(n); }
inline void mju_free(void* p) { std::free(p); }
inline mjtNum mju_abs(mjtNum x) { return std::fabs(x); }
inline mjtNum mju_pow(mjtNum x, mjtNum y) { return std::pow(x, y); }
inline mjtNum mju_exp(mjtNum x) { return std::exp(x); }
inline mjtNum mju_log(mjtNum x) { return std::log(x); }
inline mjtNum mju_log10(mjtNum x) { return std::log10(x); }
inline mjtNum mju_sqrt(mjtNum x) { return std::sqrt(x); }
inline mjtNum mju_sin(mjtNum x) { return std::sin(x); }
inline mjtNum mju_cos(mjtNum x) { return std::cos(x); }
inline mjtNum mju_tan(mjtNum x) { return std::tan(x); }
inline mjtNum mju_acos(mjtNum x) { return std::acos(x); }
inline mjtNum mju_asin(mjtNum x) { return std::asin(x); }
inline mjtNum mju_atan2(mjtNum y, mjtNum x) { return std::atan2(y, x); }
inline mjtNum mju_tanh(mjtNum x) { return std::tanh(x); }
inline mjtNum mju_floor(mjtNum x) { return std::floor(x); }
inline mjtNum mju_ceil(mjtNum x) { return std::ceil(x); }
inline mjtNum mju_round(mjtNum x) { return std::round(x); }
inline mjtNum mju_min(mjtNum a, mjtNum b) { return a <= b ? a : b; }
inline mjtNum mju_max(mjtNum a, mjtNum b) { return a >= b ? a : b; }
inline mjtNum mju_clip(mjtNum x, mjtNum lo, mjtNum hi) { return x < lo ? lo : (x > hi ? hi : x); }
inline mjtNum mju_sign(mjtNum x) { return x < 0 ? -1 : (x > 0 ? 1 : 0); }
inline int mju_isBad(mjtNum x) { return x != x || x > mjMAXVAL || x < -mjMAXVAL; }
inline void mju_zero(mjtNum* r, int n) { for (int i = 0; i < n; i++) r[i] = 0; }
inline void mju_zero3(mjtNum r[3]) { r[0] = r[1] = r[2] = 0; }
inline void mju_zero4(mjtNum r[4]) { r[0] = r[1] = r[2] = r[3] = 0; }
inline void mju_fill(mjtNum* r, mjtNum v, int n) { for (int i = 0; i < n; i++) r[i] = v; }
inline void mju_copy(mjtNum* r, const mjtNum* v, int n) { if (n > 0) std::memcpy(r, v, sizeof(mjtNum) * (size_t)n); }
inline void mju_copy3(mjtNum r[3], const mjtNum v[3]) { r[0] = v[0]; r[1] = v[1]; r[2] = v[2]; }
inline void mju_copy4(mjtNum r[4], const mjtNum v[4]) { r[0] = v[0]; r[1] = v[1]; r[2] = v[2]; r[3] = v[3]; }
inline void mju_unit4(mjtNum r[4]) { r[0] = 1; r[1] = r[2] = r[3] = 0; }
inline void mju_scl(mjtNum* r, const mjtNum* v, mjtNum s, int n) { for (int i = 0; i < n; i++) r[i] = v[i] * s; }
inline void mju_scl3(mjtNum r[3], const mjtNum v[3], mjtNum s) { r[0] = v[0] * s; r[1] = v[1] * s; r[2] = v[2] * s; }
inline void mju_add(mjtNum* r, const mjtNum* a, const mjtNum* b, int n) { for (int i = 0; i < n; i++) r[i] = a[i] + b[i]; }
inline void mju_add3(mjtNum r[3], const mjtNum a[3], const mjtNum b[3]) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
inline void mju_sub(mjtNum* r, const mjtNum* a, const mjtNum* b, int n) { for (int i = 0; i < n; i++) r[i] = a[i] - b[i]; }
inline void mju_sub3(mjtNum r[3], const mjtNum a[3], const mjtNum b[3]) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
inline void mju_addTo(mjtNum* r, const mjtNum* v, int n) { for (int i = 0; i < n; i++) r[i] += v[i]; }
inline void mju_addTo3(mjtNum r[3], const mjtNum v[3]) { r[0] += v[0]; r[1] += v[1]; r[2] += v[2]; }
inline void mju_subFrom(mjtNum* r, const mjtNum* v, int n) { for (int i = 0; i < n; i++) r[i] -= v[i]; }
inline void mju_subFrom3(mjtNum r[3], const mjtNum v[3]) { r[0] -= v[0]; r[1] -= v[1]; r[2] -= v[2]; }
inline void mju_addToScl(mjtNum* r, const mjtNum* v, mjtNum s, int n) { for (int i = 0; i < n; i++) r[i] += v[i] * s; }
inline void mju_addToScl3(mjtNum r[3], const mjtNum v[3], mjtNum s) { r[0] += v[0] * s; r[1] += v[1] * s; r[2] += v[2] * s; }
inline void mju_addScl(mjtNum* r, const mjtNum* a, const mjtNum* b, mjtNum s, int n) { for (int i = 0; i < n; i++) r[i] = a[i] + b[i] * s; }
inline void mju_addScl3(mjtNum r[3], const mjtNum a[3], const mjtNum b[3], mjtNum s) { r[0] = a[0] + b[0] * s; r[1] = a[1] + b[1] * s; r[2] = a[2] + b[2] * s; }
inline mjtNum mju_dot(const mjtNum* a, const mjtNum* b, int n) { mjtNum s = 0; for (int i = 0; i < n; i++) s += a[i] * b[i]; return s; }
inline mjtNum mju_dot3(const mjtNum a[3], const mjtNum b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline mjtNum mju_norm(const mjtNum* v, int n) { return std::sqrt(mju_dot(v, v, n)); }
inline mjtNum mju_norm3(const mjtNum v[3]) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
inline mjtNum mju_dist3(const mjtNum a[3], const mjtNum b[3]) { mjtNum d[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; return mju_norm3(d); }
inline mjtNum mju_normalize(mjtNum* v, int n) {
  mjtNum norm = std::sqrt(mju_dot(v, v, n));
  if (norm < mjMINVAL) { v[0] = 1; for (int i = 1; i < n; i++) v[i] = 0; } else { mjtNum inv = 1 / norm; for (int i = 0; i < n; i++) v[i] *= inv; }
  return norm;
}
inline mjtNum mju_normalize3(mjtNum v[3]) {
  mjtNum norm = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (norm < mjMINVAL) { v[0] = 1; v[1] = 0; v[2] = 0; } else { mjtNum inv = 1 / norm; v[0] *= inv; v[1] *= inv; v[2] *= inv; }
  return norm;
}
inline mjtNum mju_normalize4(mjtNum q[4]) {
  mjtNum norm = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (norm < mjMINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else if (std::fabs(norm - 1) > mjMINVAL) { mjtNum inv = 1 / norm; q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv; }
  return norm;
}
inline void mju_cross(mjtNum r[3], const mjtNum a[3], const mjtNum b[3]) {
  mjtNum x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0]; r[0] = x; r[1] = y; r[2] = z;
}
inline void mju_mulMatVec(mjtNum* r, const mjtNum* M, const mjtNum* v, int nr, int nc) { for (int i = 0; i < nr; i++) r[i] = mju_dot(M + i * nc, v, nc); }
inline void mju_mulMatTVec(mjtNum* r, const mjtNum* M, const mjtNum* v, int nr, int nc) {
  mju_zero(r, nc); for (int i = 0; i < nr; i++) if (v[i] != 0) mju_addToScl(r, M + i * nc, v[i], nc);
}
inline void mju_mulMatVec3(mjtNum r[3], const mjtNum M[9], const mjtNum v[3]) {
  mjtNum x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2], y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2], z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void mju_mulMatTVec3(mjtNum r[3], const mjtNum M[9], const mjtNum v[3]) {
  mjtNum x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2], y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2], z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void mju_transpose(mjtNum* r, const mjtNum* M, int nr, int nc) { for (int i = 0; i < nr; i++) for (int j = 0; j < nc; j++) r[j * nr + i] = M[i * nc + j]; }
inline void mju_mulMatMat(mjtNum* r, const mjtNum* A, const mjtNum* B, int r1, int c1, int c2) {
  mju_zero(r, r1 * c2);
  for (int i = 0; i < r1; i++) for (int k = 0; k < c1; k++) { mjtNum a = A[i * c1 + k]; if (a != 0) mju_addToScl(r + i * c2, B + k * c2, a, c2); }
}
inline void mju_mulMatMatT(mjtNum* r, const mjtNum* A, const mjtNum* B, int r1, int c1, int r2) {
  for (int i = 0; i < r1; i++) for (int j = 0; j < r2; j++) r[i * r2 + j] = mju_dot(A + i * c1, B + j * c1, c1);
}
inline void mju_mulMatTMat(mjtNum* r, const mjtNum* A, const mjtNum* B, int r1, int c1, int c2) {
  mju_zero(r, c1 * c2);
  for (int i = 0; i < r1; i++) for (int j = 0; j < c1; j++) { mjtNum a = A[i * c1 + j]; if (a != 0) mju_addToScl(r + j * c2, B + i * c2, a, c2); }
}
inline void mju_sqrMatTD(mjtNum* r, const mjtNum* M, const mjtNum* diag, int nr, int nc) {
  mju_zero(r, nc * nc);
  for (int i = 0; i < nr; i++) for (int j = 0; j < nc; j++) { mjtNum t = M[i * nc + j] * (diag ? diag[i] : 1); if (t != 0) mju_addToScl(r + j * nc, M + i * nc, t, nc); }
}
inline void mju_symmetrize(mjtNum* r, const mjtNum* M, int n) { for (int i = 0; i < n; i++) for (int j = 0; j <= i; j++) r[i * n + j] = r[j * n + i] = 0.5 * (M[i * n + j] + M[j * n + i]); }
inline void mju_eye(mjtNum* M, int n) { mju_zero(M, n * n); for (int i = 0; i < n; i++) M[i * n + i] = 1; }
inline void mju_mulQuat(mjtNum r[4], const mjtNum a[4], const mjtNum b[4]) {
  mjtNum w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  mjtNum y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
inline void mju_negQuat(mjtNum r[4], const mjtNum q[4]) { r[0] = q[0]; r[1] = -q[1]; r[2] = -q[2]; r[3] = -q[3]; }
inline void mju_quat2Mat(mjtNum m[9], const mjtNum q[4]) {
  if (q[0] == 1 && q[1] == 0 && q[2] == 0 && q[3] == 0) { m[0] = 1; m[1] = 0; m[2] = 0; m[3] = 0; m[4] = 1; m[5] = 0; m[6] = 0; m[7] = 0; m[8] = 1; return; }
  mjtNum q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3], q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  mjtNum q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2 * (q12 - q03); m[2] = 2 * (q13 + q02); m[3] = 2 * (q12 + q03); m[5] = 2 * (q23 - q01); m[6] = 2 * (q13 - q02); m[7] = 2 * (q23 + q01);
}
inline void mju_rotVecQuat(mjtNum r[3], const mjtNum v[3], const mjtNum q[4]) {
  if (v[0] == 0 && v[1] == 0 && v[2] == 0) { mju_zero3(r); return; }
  if (q[0] == 1 && q[1] == 0 && q[2] == 0 && q[3] == 0) { mju_copy3(r, v); return; }
  mjtNum m[9]; mju_quat2Mat(m, q); mju_mulMatVec3(r, m, v);
}
inline void mju_axisAngle2Quat(mjtNum r[4], const mjtNum axis[3], mjtNum angle) {
  if (angle == 0) { mju_unit4(r); return; }
  mjtNum s = std::sin(angle * 0.5); r[0] = std::cos(angle * 0.5); r[1] = axis[0] * s; r[2] = axis[1] * s; r[3] = axis[2] * s;
}
inline void mju_quat2Vel(mjtNum r[3], const mjtNum q[4], mjtNum dt) {
  mjtNum axis[3] = {q[1], q[2], q[3]};
  mjtNum sin_a_2 = mju_normalize3(axis);
  mjtNum speed = 2 * std::atan2(sin_a_2, q[0]);
  if (speed > mjPI) speed -= 2 * mjPI;
  speed /= dt;
  mju_scl3(r, axis, speed);
}
inline void mju_subQuat(mjtNum r[3], const mjtNum qa[4], const mjtNum qb[4]) {
  mjtNum qneg[4], qdif[4]; mju_negQuat(qneg, qb); mju_mulQuat(qdif, qneg, qa); mju_quat2Vel(r, qdif, 1);
}
inline void mju_quatIntegrate(mjtNum q[4], const mjtNum v[3], mjtNum s) {
  mjtNum axis[3] = {v[0], v[1], v[2]}, qr[4], t[4];
  mjtNum angle = s * mju_normalize3(axis);
  mju_axisAngle2Quat(qr, axis, angle); mju_normalize4(q); mju_mulQuat(t, q, qr); mju_copy4(q, t);
}
// dense Cholesky (engine_util_solve.c): in place, returns the rank
inline int mju_cholFactor(mjtNum* mat, int n, mjtNum mindiag) {
  int rank = n;
  for (int j = 0; j < n; j++) {
    mjtNum tmp = mat[j * (n + 1)];
    if (j) tmp -= mju_dot(mat + j * n, mat + j * n, j);
    if (tmp < mindiag) { tmp = mindiag; rank--; }
    mat[j * (n + 1)] = std::sqrt(tmp);
    tmp = 1 / mat[j * (n + 1)];
    for (int i = j + 1; i < n; i++) mat[i * n + j] = (mat[i * n + j] - mju_dot(mat + i * n, mat + j * n, j)) * tmp;
  }
  return rank;
}
inline void mju_cholSolve(mjtNum* res, const mjtNum* mat, const mjtNum* vec, int n) {
  if (res != vec) mju_copy(res, vec, n);
  for (int i = 0; i < n; i++) { if (i) res[i] -= mju_dot(mat + i * n, res, i); res[i] /= mat[i * (n + 1)]; }
  for (int i = n - 1; i >= 0; i--) { if (i < n - 1) for (int j = i + 1; j < n; j++) res[i] -= mat[j * n + i] * res[j]; res[i] /= mat[i * (n + 1)]; }
}
int mju_eig3(mjtNum eigval[3], mjtNum eigvec[9], mjtNum quat[4], const mjtNum mat[9]);
void mju_mat2Quat(mjtNum quat[4], const mjtNum mat[9]);
mjtNum mju_cholFactorBand(mjtNum* mat, int ntotal, int nband, int ndense, mjtNum diagadd, mjtNum diagmul);
void mju_cholSolveBand(mjtNum* res, const mjtNum* mat, const mjtNum* vec, int ntotal, int nband, int ndense);
void mju_band2Dense(mjtNum* res, const mjtNum* mat, int ntotal, int nband, int ndense, mjtByte flg_sym);
void mju_dense2Band(mjtNum* res, const mjtNum* mat, int ntotal, int nband, int ndense);
int mju_bandDiag(int i, int ntotal, int nband, int ndense);
inline void mju_n2f(float* r, const mjtNum* v, int n) { for (int i = 0; i < n; i++) r[i] = (float)v[i]; }
inline void mju_f2n(mjtNum* r, const float* v, int n) { for (int i = 0; i < n; i++) r[i] = v[i]; }
inline char* mju_strncpy(char* dst, const char* src, int n) { if (dst && src && n > 0) { std::strncpy(dst, src, n); dst[n - 1] = 0; } return dst; }
#endif
