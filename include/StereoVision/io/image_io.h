// B200 facade of the `.stevimg` part of LibStevi's io/image_io.h (:29-165): the container the library uses to exchange arrays
// with other LibStevi installs — one text line "<dtype> <nDim> <shape...> <strides...>" (dtype = i|u|f + bit count,
// utils/types_manipulations.h:82-102; strides in elements) followed by the elements as they lie in memory. Host-side file
// io: nothing to offload; written against Multidim::Array so fixtures made by either library load in the other (checked in
// tests/test_stevimg.py against the reference's own reader and writer). The CImg / libjpeg / libpng / libtiff formats of
// readImage / writeImage (:170-178) are outside the dense-stereo path.
#ifndef STEVI_B200_IO_IMAGE_IO_H
#define STEVI_B200_IO_IMAGE_IO_H

#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <type_traits>
#include <MultidimArrays/MultidimArrays.h>

namespace StereoVision {
namespace TypesManipulations {
// utils/types_manipulations.h:82-102
template<typename T>
inline std::string dtypeDescr() {
	if (!std::is_integral_v<T> && !std::is_floating_point_v<T>) return "";
	const char symbol = std::is_floating_point_v<T> ? 'f' : (std::is_signed_v<T> ? 'i' : 'u');
	return symbol + std::to_string(sizeof(T) * 8);
}
template<typename T>
inline bool matchdescr(std::string descr) { return descr == dtypeDescr<T>(); }
} // namespace TypesManipulations

namespace IO {

namespace B200Detail {
inline bool header(std::string const& fileName, std::string& type, int& nDim, std::stringstream& rest, std::ifstream& in) {
	in.open(fileName, std::ios_base::in | std::ios_base::binary);
	if (!in.is_open()) return false;
	std::string line;
	std::getline(in, line);
	rest.str(line);
	rest >> type >> nDim;
	return !rest.fail();
}
}

// :29-64
template<typename ImgType, int nDim>
bool stevImgFileMatchTypeAndDim(std::string const& fileName) {
	std::ifstream in; std::stringstream s; std::string type; int n = 0;
	if (!B200Detail::header(fileName, type, n, s, in)) return false;
	return type == TypesManipulations::dtypeDescr<ImgType>() && n <= nDim;
}

// :66-127: converts to ImgType, makes the array dense, writes header + elements
template<typename ImgType, typename InType, int nDim>
bool writeStevimg(std::string const& fileName, Multidim::Array<InType, nDim> const& image) {
	if constexpr (!std::is_same_v<ImgType, InType>) {
		return writeStevimg<ImgType, ImgType, nDim>(fileName, image.template cast<ImgType>());
	} else {
		if (!image.isDense() && !image.empty()) {
			Multidim::Array<InType, nDim> dense(image.shape());
			typename Multidim::Array<InType, nDim>::IndexBlock idx;
			idx.setZero();
			for (int k = 0; k < image.flatLenght(); k++) { dense.atUnchecked(idx) = image.valueUnchecked(idx); idx.moveToNextIndex(image.shape()); }
			return writeStevimg<ImgType, InType, nDim>(fileName, dense);
		}
		std::FILE* f = std::fopen(fileName.c_str(), "wb");
		if (!f) return false;
		std::string head = TypesManipulations::dtypeDescr<ImgType>() + ' ' + std::to_string(nDim);
		for (int k = 0; k < nDim; k++) head += ' ' + std::to_string(image.shape()[k]);
		for (int k = 0; k < nDim; k++) head += ' ' + std::to_string(image.strides()[k]);
		head += '\n';
		bool ok = std::fwrite(head.data(), 1, head.size(), f) == head.size();
		const size_t n = size_t(image.flatLenght());
		if (n > 0) ok = ok && std::fwrite(&const_cast<Multidim::Array<InType, nDim>&>(image).atUnchecked(0), sizeof(ImgType), n, f) == n;
		return (std::fclose(f) == 0) && ok;
	}
}

// :129-165: empty array when the file is missing, holds another type, or has more dimensions; missing dimensions are 1
template<typename ImgType, int nDim>
Multidim::Array<ImgType, nDim> readStevimg(std::string const& fileName) {
	std::ifstream in; std::stringstream s; std::string type; int n = 0;
	if (!B200Detail::header(fileName, type, n, s, in)) return Multidim::Array<ImgType, nDim>();
	if (type != TypesManipulations::dtypeDescr<ImgType>() || n > nDim) return Multidim::Array<ImgType, nDim>();
	typename Multidim::Array<ImgType, nDim>::ShapeBlock shape, stride;
	for (int k = 0; k < nDim; k++) { if (k < n) s >> shape[k]; else shape[k] = 1; }
	for (int k = 0; k < nDim; k++) { if (k < n) s >> stride[k]; else stride[k] = 1; }
	Multidim::Array<ImgType, nDim> img(shape, stride);
	if (img.flatLenght() > 0) in.read(reinterpret_cast<char*>(&img.atUnchecked(0)), std::streamsize(sizeof(ImgType)) * img.flatLenght());
	return img;
}

} // namespace IO
} // namespace StereoVision

#endif
