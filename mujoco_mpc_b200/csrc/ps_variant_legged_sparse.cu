// Engine instantiation "legged_sparse": thread-per-rollout, compressed-Jacobian phases, quadruped capacities.
#include "ps_variant.h"
namespace ps {
const Variant* variant_legged_sparse() { static const Variant v = make_sparse_variant<LeggedCoopT, 64>("legged_sparse"); return &v; }
}  // namespace ps
