// multigpu.cu — in-process multi-GPU render behind the launch boundary (`numGpus` of novum_launch_*): one host thread per GPU.
#include "scene.h"

namespace nv {

int launchMultiGpu(int numGpus, int integrator, const NovumRenderConfig& cfg, const NovumCamera& cam, NovumScene* primary, int numSample,
                   unsigned long long seed, float4* colors, int useMIS) {
    (void)numGpus; (void)integrator; (void)cfg; (void)cam; (void)primary; (void)numSample; (void)seed; (void)colors; (void)useMIS;
    return NOVUM_E_UNSUPPORTED;
}

}  // namespace nv
