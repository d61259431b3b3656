// Engine instantiation "tiny" (capacities: ps_host_model.h TinyT).
#define PS_THREAD_INLINE 1
#include "ps_variant.h"
namespace ps {
const Variant* variant_tiny() { static const Variant v = make_variant<TinyT, 128>("tiny"); return &v; }
}  // namespace ps
