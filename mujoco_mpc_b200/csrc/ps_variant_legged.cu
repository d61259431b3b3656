// Engine instantiation "legged" (capacities: ps_host_model.h LeggedT).
#include "ps_variant.h"
namespace ps {
const Variant* variant_legged() { static const Variant v = [] { Variant x = make_variant<LeggedT, 64>("legged", launch_robust_legged); x.fixup = launch_fixup_legged; return x; }(); return &v; }
}  // namespace ps
