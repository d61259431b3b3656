// ORACLE (test infrastructure).  Quadruped residual (mjpc/tasks/quadruped/quadruped.cc:33-221).
// Placeholder until the quadruped row of SURVEY.md §8 is built: writes zeros.
#include "mm_planner.h"
namespace mm {
void quadrupedResidual(const Model* m, const Data* d, const PsTask* t, double* r) {
  (void)m; (void)d; for (int i = 0; i < t->num_residual; i++) r[i] = 0;
}
}  // namespace mm
