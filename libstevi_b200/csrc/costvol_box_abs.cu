// Instantiations of the box-sum cost-volume kernel (costvol_box.cuh) for one pair operation; split per file so the
// instantiations compile in parallel.
#include "costvol_box.cuh"

namespace stevi {
int32_t launch_cv_box_abs(stevi_ctx* c, const BoxParams& p, int hr, int sign) { return launch_cv_box<BOX_ABS>(c, p, hr, sign); }
} // namespace stevi
