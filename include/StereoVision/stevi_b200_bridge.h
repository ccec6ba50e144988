// Bridge between the reference's C++ template surface (namespace StereoVision::Correlation over Multidim::Array) and
// the C ABI of the B200 library (include/stevi_b200/c_api.h). Header-only, C++17, no CUDA or torch types.
//
// The headers next to this one re-create the names, template parameter lists and defaults of the LibStevi headers of the
// same name for the dense-stereo hot path; their bodies describe the operands, allocate the result
// with the shape and strides the reference would give it, and forward to the C ABI. Any non-zero status maps to the
// reference's error convention: an empty array. Instantiations the library does not cover fail to compile
// (static_assert) instead of silently computing on the CPU — there is no CPU fallback.
#ifndef STEVI_B200_BRIDGE_H
#define STEVI_B200_BRIDGE_H

#include <cstdint>
#include <type_traits>
#include <array>
#include <MultidimArrays/MultidimArrays.h>
#include <stevi_b200/c_api.h>

namespace StereoVision {
namespace B200 {

template<class T> constexpr int dtypeOf() {
	static_assert(std::is_same_v<T, uint8_t> || std::is_same_v<T, uint16_t> || std::is_same_v<T, int32_t> || std::is_same_v<T, int> ||
				  std::is_same_v<T, uint32_t> || std::is_same_v<T, float>, "element type not covered by the B200 library");
	return std::is_same_v<T, uint8_t> ? STEVI_U8 : std::is_same_v<T, uint16_t> ? STEVI_U16 :
		   (std::is_same_v<T, int32_t> || std::is_same_v<T, int>) ? STEVI_I32 : std::is_same_v<T, uint32_t> ? STEVI_U32 : STEVI_F32;
}

// one context per host thread: the reference's functions are stateless and re-entrant, so is this
inline stevi_ctx* threadContext() {
	struct Holder { stevi_ctx* c = nullptr; ~Holder() { if (c) stevi_ctx_destroy(c); } };
	thread_local Holder h;
	if (!h.c) stevi_ctx_create(0, nullptr, &h.c);
	return h.c;   // nullptr when there is no CUDA device: every call then fails and the facade returns empty arrays
}

// descriptor of a host array (a view: nothing is copied)
template<class T, int n, Multidim::ArrayDataAccessConstness C>
stevi_array describe(Multidim::Array<T, n, C> const& a) {
	static_assert(n >= 1 && n <= STEVI_MAX_DIMS, "rank");
	stevi_array d{};
	typename Multidim::Array<T, n, C>::IndexBlock zero;
	zero.fill(0);
	d.data = a.empty() ? nullptr : const_cast<T*>(&const_cast<Multidim::Array<T, n, C>&>(a).atUnchecked(zero));
	d.dtype = dtypeOf<T>();
	d.ndim = n;
	d.mem = STEVI_HOST;
	auto sh = a.shape();
	auto st = a.strides();
	for (int k = 0; k < n; k++) { d.shape[k] = sh[k]; d.strides[k] = st[k]; }
	return d;
}

inline bool ok(int32_t status) { return status == STEVI_OK; }

} // namespace B200
} // namespace StereoVision

#endif
