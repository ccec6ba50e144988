// Stand-in for french-paragon/MultidimArrays' MultidimIndexManipulators.h (un-vendored dependency of the reference, absent
// from this image). Test infrastructure: it only has to let interpolation/interpolation.h parse when
// correlation/image_based_refinement.h is compiled into oracle/_ref; the two classes below are functional in the plain
// row-major sense (last index fastest), which nothing built here relies on.
#ifndef STEVI_ORACLE_MULTIDIM_INDEX_MANIPULATORS_STUB
#define STEVI_ORACLE_MULTIDIM_INDEX_MANIPULATORS_STUB
#include <array>
#include <MultidimArrays/MultidimArrays.h>
namespace Multidim {
template<int N> class DimsExclusionSet {
public:
	template<typename... Ds> DimsExclusionSet(Ds... ds) : _n(0) { int v[] = {int(ds)..., 0}; for (size_t k = 0; k < sizeof...(Ds); k++) _ex[_n++] = v[k]; }
	bool excluded(int d) const { for (int k = 0; k < _n; k++) if (_ex[k] == d) return true; return false; }
private:
	int _ex[N + 1]; int _n;
};
template<int N> class IndexConverter {
public:
	template<typename S> explicit IndexConverter(S const& shape) { for (int k = 0; k < N; k++) { _s[k] = shape[k]; _on[k] = true; } }
	template<typename S> IndexConverter(S const& shape, DimsExclusionSet<N> const& ex) { for (int k = 0; k < N; k++) { _s[k] = shape[k]; _on[k] = !ex.excluded(k); } }
	int numberOfPossibleIndices() const { int n = 1; for (int k = 0; k < N; k++) if (_on[k]) n *= _s[k]; return n; }
	std::array<int, N> getIndexFromPseudoFlatId(int id) const {
		std::array<int, N> r{};
		for (int k = N - 1; k >= 0; k--) { if (!_on[k]) { r[k] = 0; continue; } r[k] = id % _s[k]; id /= _s[k]; }
		return r;
	}
private:
	int _s[N]; bool _on[N];
};
template<size_t N> IndexConverter(std::array<int, N> const&) -> IndexConverter<int(N)>;
template<size_t N> IndexConverter(std::array<int, N> const&, DimsExclusionSet<int(N)> const&) -> IndexConverter<int(N)>;
}
#endif
