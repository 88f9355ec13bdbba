// placeholder so that sources including <faiss/Heap.h> through the ivf-hnsw headers keep compiling;
// the drop-in needs nothing from it.
#pragma once
