// placeholder so that sources including <faiss/FaissAssert.h> through the ivf-hnsw headers keep compiling;
// the drop-in needs nothing from it.
#pragma once
