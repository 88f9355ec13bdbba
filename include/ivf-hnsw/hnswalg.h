// <ivf-hnsw/hnswalg.h> as installed by the reference's setup_env_dev.sh (symlink to hnswlib/hnswalg.h)
#pragma once
#include <hnswlib/hnswalg.h>
