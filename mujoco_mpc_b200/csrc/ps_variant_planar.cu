// Engine instantiation "planar" (capacities: ps_host_model.h PlanarT).
#define PS_THREAD_INLINE 1
#include "ps_variant.h"
namespace ps {
const Variant* variant_planar() { static const Variant v = make_variant<PlanarT, 64>("planar", launch_robust_planar); return &v; }
}  // namespace ps
