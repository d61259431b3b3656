// Engine instantiation "planar_coop": warp-per-rollout engine at the planar capacities (walker).
#include "ps_variant.h"
namespace ps {
const Variant* variant_planar_coop() { static const Variant v = make_coop_variant<PlanarCoopT, 4>("planar_coop", launch_robust_planar_coop); return &v; }
}  // namespace ps
