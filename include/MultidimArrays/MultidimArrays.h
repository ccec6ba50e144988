// Minimal stand-in for french-paragon/MultidimArrays (pinned by the reference at
// 78d674a3a3713feee4ced37a95fef3b53cc2832b, /root/reference/CMakeLists.txt:62-67), which is
// fetched at configure time upstream and is NOT vendored in the reference tree.
//
// Only the API subset used on the dense-stereo hot path is provided.
// It does indexing only (offset = sum idx_i * stride_i), no arithmetic. Two behaviours leak into
// results and are fixed here by fiat (they are not pinned by any reference test):
//   * allocations are zero-initialised (makes the never-written census word deterministic, Q1);
//   * the default stride order is first-index-fastest unless MULTIDIM_DEFAULT_C_ORDER is defined.
//
// Used both by the oracle build (reference headers compiled unmodified against it) and by the
// drop-in facade under include/StereoVision/.
#ifndef STEVI_B200_MULTIDIMARRAYS_SHIM_H
#define STEVI_B200_MULTIDIMARRAYS_SHIM_H

#include <array>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <type_traits>
#include <initializer_list>
#include <algorithm>
#include <cmath>
#include <vector>

namespace Multidim {

typedef int array_size_t;

enum class AccessCheck { Check, Nocheck };
enum ArrayDataAccessConstness { NonConstView = 0, ConstView = 1 };

struct DimIndex {
	explicit DimIndex(int i) : index(i) {}
	int index;
};

struct DimSlice {
	DimSlice() : begin(0), end(-1), whole(true) {}
	DimSlice(int b, int e) : begin(b), end(e), whole(false) {}
	int begin;
	int end; // half-open
	bool whole;
};

template<int nDim>
struct IndexBlockT : public std::array<int, nDim> {
	IndexBlockT() { this->fill(0); }
	IndexBlockT(std::array<int, nDim> const& a) : std::array<int, nDim>(a) {}
	IndexBlockT(std::initializer_list<int> l) {
		this->fill(0);
		int k = 0;
		for (int v : l) { if (k < nDim) (*this)[k++] = v; }
	}
	void setZero() { this->fill(0); }
	IndexBlockT operator+(IndexBlockT const& o) const { IndexBlockT r; for (int k = 0; k < nDim; k++) r[k] = (*this)[k] + o[k]; return r; }
	IndexBlockT operator-(IndexBlockT const& o) const { IndexBlockT r; for (int k = 0; k < nDim; k++) r[k] = (*this)[k] - o[k]; return r; }
	IndexBlockT operator+(int v) const { IndexBlockT r; for (int k = 0; k < nDim; k++) r[k] = (*this)[k] + v; return r; }
	IndexBlockT operator-(int v) const { IndexBlockT r; for (int k = 0; k < nDim; k++) r[k] = (*this)[k] - v; return r; }
	bool isInLimit(std::array<int, nDim> const& shape) const { for (int k = 0; k < nDim; k++) if ((*this)[k] < 0 || (*this)[k] >= shape[k]) return false; return true; }
	void clip(std::array<int, nDim> const& shape) { for (int k = 0; k < nDim; k++) (*this)[k] = (*this)[k] < 0 ? 0 : ((*this)[k] >= shape[k] ? shape[k] - 1 : (*this)[k]); }
	// first index fastest, like upstream's flat traversal
	void moveToNextIndex(std::array<int, nDim> const& shape) {
		for (int k = 0; k < nDim; k++) {
			(*this)[k]++;
			if ((*this)[k] < shape[k]) return;
			(*this)[k] = 0;
		}
	}
};

template<typename T, int nDim, ArrayDataAccessConstness viewConstness = NonConstView>
class Array {
public:
	typedef std::array<int, nDim> ShapeBlock;
	typedef IndexBlockT<nDim> IndexBlock;
	typedef T ScalarT;
	static constexpr int nDims = nDim;

	template<typename... Ds>
	static constexpr int countSlices() {
		return (0 + ... + (std::is_same<typename std::decay<Ds>::type, DimSlice>::value ? 1 : 0));
	}

	static ShapeBlock defaultStrides(ShapeBlock const& shape) {
		ShapeBlock st;
		long s = 1;
#ifdef MULTIDIM_DEFAULT_C_ORDER
		for (int k = nDim - 1; k >= 0; k--) { st[k] = int(s); s *= shape[k]; }
#else
		for (int k = 0; k < nDim; k++) { st[k] = int(s); s *= shape[k]; }
#endif
		return st;
	}

	Array() : _data(nullptr), _manage(false) { _shape.fill(0); _strides.fill(0); }

	template<typename... Ds,
			 typename = typename std::enable_if<(sizeof...(Ds) == nDim) &&
				 (std::is_integral<typename std::decay<Ds>::type>::value && ...)>::type>
	explicit Array(Ds... dims) : _manage(true) {
		_shape = ShapeBlock{static_cast<int>(dims)...};
		_strides = defaultStrides(_shape);
		allocate();
	}

	explicit Array(ShapeBlock const& shape) : _shape(shape), _manage(true) {
		_strides = defaultStrides(_shape);
		allocate();
	}

	Array(ShapeBlock const& shape, ShapeBlock const& strides) : _shape(shape), _strides(strides), _manage(true) {
		allocate();
	}

	Array(T* data, ShapeBlock const& shape, ShapeBlock const& strides, bool manage = false)
		: _data(data), _shape(shape), _strides(strides), _manage(manage) {}

	// copy: deep when owning, alias when a view
	Array(Array const& o) : _shape(o._shape), _strides(o._strides), _manage(o._manage) {
		if (o._manage && o._data != nullptr) {
			size_t n = o.span();
			_data = new T[n];
			std::copy(o._data, o._data + n, _data);
		} else {
			_data = o._data;
		}
	}

	Array(Array&& o) noexcept : _data(o._data), _shape(o._shape), _strides(o._strides), _manage(o._manage) {
		o._data = nullptr;
		o._manage = false;
	}

	// constness converting constructors (alias; ownership moves for rvalues)
	template<ArrayDataAccessConstness oc, typename = typename std::enable_if<oc != viewConstness>::type>
	Array(Array<T, nDim, oc> const& o)
		: _data(const_cast<T*>(o.rawData())), _shape(o.shape()), _strides(o.strides()), _manage(false) {}

	template<ArrayDataAccessConstness oc, typename = typename std::enable_if<oc != viewConstness>::type>
	Array(Array<T, nDim, oc>&& o)
		: _data(const_cast<T*>(o.rawData())), _shape(o.shape()), _strides(o.strides()), _manage(o.managesData()) {
		o.releaseOwnership();
	}

	~Array() { if (_manage && _data != nullptr) delete[] _data; }

	Array& operator=(Array const& o) {
		if (this == &o) return *this;
		Array tmp(o);
		swap(tmp);
		return *this;
	}
	Array& operator=(Array&& o) noexcept {
		if (this == &o) return *this;
		if (_manage && _data != nullptr) delete[] _data;
		_data = o._data; _shape = o._shape; _strides = o._strides; _manage = o._manage;
		o._data = nullptr; o._manage = false;
		return *this;
	}

	void swap(Array& o) {
		std::swap(_data, o._data); std::swap(_shape, o._shape);
		std::swap(_strides, o._strides); std::swap(_manage, o._manage);
	}

	ShapeBlock shape() const { return _shape; }
	ShapeBlock strides() const { return _strides; }
	bool empty() const {
		if (_data == nullptr) return true;
		for (int k = 0; k < nDim; k++) if (_shape[k] <= 0) return true;
		return false;
	}
	int flatLenght() const { // (sic) upstream spelling
		long n = 1;
		for (int k = 0; k < nDim; k++) n *= _shape[k];
		return int(n);
	}
	bool isDense() const { return span() == size_t(std::max(flatLenght(), 0)); }
	bool managesData() const { return _manage; }
	void releaseOwnership() { _manage = false; }
	T const* rawData() const { return _data; }
	T* takePointer() { _manage = false; return _data; }
	T* data() { return _data; }
	T const* data() const { return _data; }

	// number of elements from the first to one past the last addressed (non-negative strides)
	size_t span() const {
		size_t s = 1;
		for (int k = 0; k < nDim; k++) {
			if (_shape[k] <= 0) return 0;
			s += size_t(_shape[k] - 1) * size_t(std::abs(_strides[k]));
		}
		return s;
	}

	bool inBound(ShapeBlock const& idx) const {
		for (int k = 0; k < nDim; k++) if (idx[k] < 0 || idx[k] >= _shape[k]) return false;
		return true;
	}

	long offset(ShapeBlock const& idx) const {
		long o = 0;
		for (int k = 0; k < nDim; k++) o += long(idx[k]) * long(_strides[k]);
		return o;
	}

	template<AccessCheck c = AccessCheck::Check, typename... Is,
			 typename = typename std::enable_if<(sizeof...(Is) == nDim) &&
				 (std::is_integral<typename std::decay<Is>::type>::value && ...)>::type>
	T& at(Is... is) {
		ShapeBlock idx{static_cast<int>(is)...};
		if (c == AccessCheck::Check && !inBound(idx)) throw std::out_of_range("Multidim::Array::at");
		return _data[offset(idx)];
	}
	template<AccessCheck c = AccessCheck::Check>
	T& at(ShapeBlock const& idx) {
		if (c == AccessCheck::Check && !inBound(idx)) throw std::out_of_range("Multidim::Array::at");
		return _data[offset(idx)];
	}
	template<typename... Is,
			 typename = typename std::enable_if<(sizeof...(Is) == nDim) &&
				 (std::is_integral<typename std::decay<Is>::type>::value && ...)>::type>
	T& atUnchecked(Is... is) { return _data[offset(ShapeBlock{static_cast<int>(is)...})]; }
	T& atUnchecked(ShapeBlock const& idx) { return _data[offset(idx)]; }
	// flat position in the array's own memory (what io/image_io.h:96,160 use to reach the first element of a dense array)
	template<int D = nDim, typename = typename std::enable_if<(D > 1)>::type>
	T& atUnchecked(int flatPosition) { return _data[flatPosition]; }

	template<AccessCheck c = AccessCheck::Check, typename... Is,
			 typename = typename std::enable_if<(sizeof...(Is) == nDim) &&
				 (std::is_integral<typename std::decay<Is>::type>::value && ...)>::type>
	T value(Is... is) const {
		ShapeBlock idx{static_cast<int>(is)...};
		if (c == AccessCheck::Check && !inBound(idx)) throw std::out_of_range("Multidim::Array::value");
		return _data[offset(idx)];
	}
	template<AccessCheck c = AccessCheck::Check>
	T value(ShapeBlock const& idx) const {
		if (c == AccessCheck::Check && !inBound(idx)) throw std::out_of_range("Multidim::Array::value");
		return _data[offset(idx)];
	}
	template<typename... Is,
			 typename = typename std::enable_if<(sizeof...(Is) == nDim) &&
				 (std::is_integral<typename std::decay<Is>::type>::value && ...)>::type>
	T valueUnchecked(Is... is) const { return _data[offset(ShapeBlock{static_cast<int>(is)...})]; }
	T valueUnchecked(ShapeBlock const& idx) const { return _data[offset(idx)]; }

	T valueOrAlt(ShapeBlock const& idx, T const& alt) const {
		if (!inBound(idx)) return alt;
		return _data[offset(idx)];
	}

	template<typename T2>
	Array<T2, nDim> cast() const {
		Array<T2, nDim> out(_shape, _strides);
		size_t n = span();
		T2* o = out.data();
		for (size_t k = 0; k < n; k++) o[k] = static_cast<T2>(_data[k]);
		return out;
	}

	template<int nOut>
	Array<T, nOut, viewConstness> buildReshapedView(std::array<int, nOut> const& shape,
													 std::array<int, nOut> const& strides) const {
		return Array<T, nOut, viewConstness>(_data, shape, strides, false);
	}

	void copyData(Array const& o) {
		IndexBlock idx;
		int n = flatLenght();
		for (int k = 0; k < n; k++) { atUnchecked(idx) = o.valueUnchecked(idx); idx.moveToNextIndex(_shape); }
	}

	// subView: any mix of DimIndex / DimSlice, one per dimension
	template<typename... Ds, typename = typename std::enable_if<(sizeof...(Ds) == nDim)>::type>
	Array<T, countSlices<Ds...>(), viewConstness> subView(Ds... ds) const {
		constexpr int nOut = countSlices<Ds...>();
		std::array<int, nOut> shape{};
		std::array<int, nOut> strides{};
		long off = 0;
		int dim = 0;
		int odim = 0;
		auto handle = [&](auto const& d) {
			using D = typename std::decay<decltype(d)>::type;
			if constexpr (std::is_same<D, DimIndex>::value) {
				off += long(d.index) * _strides[dim];
			} else {
				int b = d.whole ? 0 : d.begin;
				int e = d.whole ? _shape[dim] : d.end;
				off += long(b) * _strides[dim];
				shape[odim] = e - b;
				strides[odim] = _strides[dim];
				odim++;
			}
			dim++;
		};
		(handle(ds), ...);
		return Array<T, nOut, viewConstness>(_data + off, shape, strides, false);
	}

	Array<T, nDim - 1, viewConstness> sliceView(int dim, int index) const {
		std::array<int, nDim - 1> shape{};
		std::array<int, nDim - 1> strides{};
		int o = 0;
		for (int k = 0; k < nDim; k++) {
			if (k == dim) continue;
			shape[o] = _shape[k]; strides[o] = _strides[k]; o++;
		}
		return Array<T, nDim - 1, viewConstness>(_data + long(index) * _strides[dim], shape, strides, false);
	}

	// the 1-D view along dimension `dim` at the position given for the other dimensions, in their order (on-demand cost volumes
	// read one pixel's feature vector with it: correlation/on_demand_cost_volume.h:164-172)
	template<typename P>
	Array<T, 1, ConstView> indexDimView(int dim, P const& otherIdxs) const {
		long off = 0;
		int o = 0;
		for (int k = 0; k < nDim; k++) {
			if (k == dim) continue;
			off += long(otherIdxs[o]) * _strides[k]; o++;
		}
		return Array<T, 1, ConstView>(_data + off, std::array<int, 1>{_shape[dim]}, std::array<int, 1>{_strides[dim]}, false);
	}

private:
	void allocate() {
		size_t n = span();
		_data = new T[n > 0 ? n : 1](); // zero-initialised, see header note
	}

	T* _data;
	ShapeBlock _shape;
	ShapeBlock _strides;
	bool _manage;
};

} // namespace Multidim

#endif // STEVI_B200_MULTIDIMARRAYS_SHIM_H
