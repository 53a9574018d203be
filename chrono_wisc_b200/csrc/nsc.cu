// nsc.cu -- NSC DVI solve (APGD + matrix-free Schur product). Placeholder until the SMC path is measured.
#include "common.cuh"
namespace cb200 {
int enqueue_nsc(System* S) {
    (void)S;
    return set_error(CB200_ERR_UNSUPPORTED, "NSC solver not built yet");
}
}  // namespace cb200
extern "C" int cb200_download_gamma(cb200_system*, double*, int) {
    return cb200::set_error(CB200_ERR_UNSUPPORTED, "NSC solver not built yet");
}
extern "C" int cb200_pack_slab_dev(cb200_system*, int, double, double, void*, int64_t, int64_t*) {
    return cb200::set_error(CB200_ERR_UNSUPPORTED, "slab exchange not built yet");
}
extern "C" int cb200_unpack_slab_dev(cb200_system*, const void*, int64_t) {
    return cb200::set_error(CB200_ERR_UNSUPPORTED, "slab exchange not built yet");
}
