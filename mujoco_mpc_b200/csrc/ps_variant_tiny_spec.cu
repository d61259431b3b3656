// Model-specialised instantiations of the lean thread-per-rollout kernel: one per registered tiny task
// (csrc/ps_spec_generated.h).  Topology tables, joint / actuator / sensor types and the cost-term layout are compile-time
// constants here, the loops over bodies / joints / dofs / terms are fully unrolled (PS_SPEC_BUILD -> PS_SPEC_UNROLL), so
// the rollout's working set lives in registers instead of local memory.  Same arithmetic, in the same order, as the
// generic kernel: results are bit-identical.
#define PS_THREAD_INLINE 1
#define PS_SPEC_BUILD 1
#define PS_ROLLOUT_SINGLE_SITE 1
#include "ps_variant.h"
#ifndef PS_SPEC_BLOCK
#define PS_SPEC_BLOCK 256
#endif
namespace ps {
#define PS_DEF_SPEC_VARIANT(N) \
  const Variant* variant_spec_##N() { static const Variant v = make_spec_variant<SpecTraits<Spec##N>, PS_SPEC_BLOCK>("tiny_" #N); return &v; }
PS_SPEC_LIST(PS_DEF_SPEC_VARIANT)
}  // namespace ps
