// Engine instantiation "planar_sparse": thread-per-rollout, compressed-Jacobian phases, planar capacities.
#include "ps_variant.h"
namespace ps {
const Variant* variant_planar_sparse() { static const Variant v = make_sparse_variant<PlanarCoopT, 64>("planar_sparse"); return &v; }
}  // namespace ps
