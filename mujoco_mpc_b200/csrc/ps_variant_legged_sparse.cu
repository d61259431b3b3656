// Engine instantiation "legged_sparse": thread-per-rollout, compressed-Jacobian phases, quadruped capacities.
#include "ps_variant.h"
#ifndef PS_SPARSE_BLOCK
#define PS_SPARSE_BLOCK 512
#endif
namespace ps {
const Variant* variant_legged_sparse() { static const Variant v = [] { Variant x = make_sparse_variant<LeggedCoopT, PS_SPARSE_BLOCK>("legged_sparse"); x.fixup = launch_fixup_legged; return x; }(); return &v; }
}  // namespace ps
