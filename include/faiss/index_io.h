// faiss file formats used by ivf-hnsw (faiss/index_io.cpp:425-513), byte compatible with the reference.
#pragma once
#include <faiss/ProductQuantizer.h>
#include <faiss/VectorTransform.h>

namespace faiss {

void write_ProductQuantizer(const ProductQuantizer *pq, const char *fname);
ProductQuantizer *read_ProductQuantizer(const char *fname);
void write_VectorTransform(const VectorTransform *vt, const char *fname);
VectorTransform *read_VectorTransform(const char *fname);

}  // namespace faiss
