// placeholder until the tcgen05 backward kernels land (same commit series): rejects every call loudly.
#include "host_common.h"
extern "C" {
int owl_attn_bwd(const void*, const void*, int64_t, int64_t, const void*, int64_t, int64_t, const void*, int64_t,
                 int64_t, const void*, int64_t, int64_t, const float*, float*, float*, void*, int64_t, int64_t, void*,
                 int64_t, int64_t, void*, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t,
                 const owl_attn_mask*, owl_stream_t) { return OWL_ERR_UNSUPPORTED; }
}
