// tcgen05 / TMEM / TMA attention kernel (placeholder until the kernel lands).
#include "tmx_attn.cuh"
namespace tmx {
int launch_attn_sm100(const AttnParams&, cudaStream_t) { return TMX_ERR_UNSUPPORTED; }
}  // namespace tmx
