// bsg_reset.cu -- device-side scenario generation (the reset path of the env layer).
#include "bsg_handle.h"

extern "C" int bsg_reset(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream) {
    (void)mask_dev; (void)seed; (void)stream;
    if (!h) return BSG_ERR_ARG;
    h->err = "bsg_reset: not implemented yet";
    return BSG_ERR_UNSUPPORTED;
}
