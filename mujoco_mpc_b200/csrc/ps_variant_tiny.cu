// Engine instantiation "tiny" (capacities: ps_host_model.h TinyT).
#define PS_THREAD_INLINE 1
#include "ps_variant.h"
#ifndef PS_TINY_BLOCK
#define PS_TINY_BLOCK 128
#endif
namespace ps {
const Variant* variant_tiny() { static const Variant v = make_variant<TinyT, PS_TINY_BLOCK>("tiny", launch_robust_tiny); return &v; }
}  // namespace ps
