// Engine instantiation "tiny_coop": warp-per-rollout engine at the tiny capacities (cartpole, acrobot, particle).
#include "ps_variant.h"
namespace ps {
const Variant* variant_tiny_coop() { static const Variant v = make_coop_variant<TinyCoopT, 8>("tiny_coop"); return &v; }
}  // namespace ps
