#include "resample_impl.cuh"
namespace txb {
TXB_RESAMPLE_FT_IMPL(int32_t)
}
