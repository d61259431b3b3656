// Engine instantiation "legged_coop": warp-per-rollout engine at the quadruped capacities.
#include "ps_variant.h"
namespace ps {
const Variant* variant_legged_coop() { static const Variant v = [] { Variant x = make_coop_variant<LeggedCoopT, 4>("legged_coop", launch_robust_legged_coop); x.fixup = launch_fixup_legged; return x; }(); return &v; }
}  // namespace ps
