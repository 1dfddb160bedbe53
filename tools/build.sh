#!/bin/bash
# rebuild libvo_b200.so in-tree (nvcc -gencode arch=compute_100a,code=sm_100a) from anywhere
cd "$(dirname "$0")/.." && python -c "
import importlib
importlib.import_module('locally-consistent-visual-odometry-pipeline_b200._build').build()
"
