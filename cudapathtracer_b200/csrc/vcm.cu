// vcm.cu — placeholder until the VCM / SPPM / BDPT engine lands (next milestone).
#include "scene.h"
namespace nv {
int vcmRenderDevice(NovumScene*, const NovumRenderConfig*, const NovumCamera*, int, int, NovumShard, unsigned long long, float4*,
                    NovumRenderStats*) { return NOVUM_E_UNSUPPORTED; }
}
extern "C" {
int novum_vcm_create(NovumScene*, const NovumRenderConfig*, const NovumCamera*, NovumShard, uint64_t, NovumVcmState**) { return NOVUM_E_UNSUPPORTED; }
void novum_vcm_destroy(NovumVcmState*) {}
int novum_vcm_light_pass(NovumVcmState*, int, NovumFloat4*) { return NOVUM_E_UNSUPPORTED; }
int novum_vcm_cache_buffer(NovumVcmState*, void**, uint64_t*, uint64_t*) { return NOVUM_E_UNSUPPORTED; }
int novum_vcm_build_grid(NovumVcmState*, int) { return NOVUM_E_UNSUPPORTED; }
int novum_vcm_eye_pass(NovumVcmState*, int, NovumFloat4*) { return NOVUM_E_UNSUPPORTED; }
int novum_vcm_stats(NovumVcmState*, NovumRenderStats*) { return NOVUM_E_UNSUPPORTED; }
}
