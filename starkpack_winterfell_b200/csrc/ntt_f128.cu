// f128 instantiations of the NTT pass kernel.
#include "ntt_launch.cuh"
namespace wg {
WG_DEFINE_LAUNCH_NTT_PASS(F128)
}
