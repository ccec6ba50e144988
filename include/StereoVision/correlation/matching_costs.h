// Vocabulary of LibStevi's correlation/matching_costs.h used on the dense-stereo path: matchingFunctions (:38-53),
// MatchingFunctionTraits (:419-685), hamming_cv_t (:234), census_data_t (:773). The comparison kernels themselves run on
// the GPU (libstevi_b200/csrc/costvol*.cu, census.cu).
#ifndef STEVI_B200_CORRELATION_MATCHING_COSTS_H
#define STEVI_B200_CORRELATION_MATCHING_COSTS_H

#include <cstdint>
#include "./correlation_base.h"

namespace StereoVision {
namespace Correlation {

enum class matchingFunctions {
	None = -1, CC = 0, NCC = 1, SSD = 2, SAD = 3, ZCC = 4, ZNCC = 5, ZSSD = 6, ZSAD = 7, MEDAD = 8, ZMEDAD = 9,
	HAMMING = 10, CENSUS = 11, KERMI = 12
};

typedef uint16_t hamming_cv_t;
typedef uint32_t census_data_t;

template<matchingFunctions func>
class MatchingFunctionTraits {
	static constexpr bool is(matchingFunctions a) { return func == a; }
public:
	static constexpr bool ZeroMean = is(matchingFunctions::ZCC) || is(matchingFunctions::ZNCC) || is(matchingFunctions::ZSSD) ||
									 is(matchingFunctions::ZSAD) || is(matchingFunctions::ZMEDAD);
	static constexpr bool Normalized = is(matchingFunctions::NCC) || is(matchingFunctions::ZNCC);
	static constexpr bool isCensusBased = is(matchingFunctions::CENSUS);
	static constexpr dispExtractionStartegy extractionStrategy =
			(is(matchingFunctions::CC) || is(matchingFunctions::NCC) || is(matchingFunctions::ZCC) || is(matchingFunctions::ZNCC))
			? dispExtractionStartegy::Score : dispExtractionStartegy::Cost;
	// instantiations the B200 library covers
	static constexpr bool onB200 = !is(matchingFunctions::None) && !is(matchingFunctions::MEDAD) && !is(matchingFunctions::ZMEDAD) &&
								   !is(matchingFunctions::KERMI) && !is(matchingFunctions::HAMMING);
};

} // namespace Correlation
} // namespace StereoVision

#endif
