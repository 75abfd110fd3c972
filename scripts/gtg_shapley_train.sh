#!/bin/sh
# the GTG-Shapley paper workload (reference: gtg_shapley_train.sh)
cd "$(dirname "$0")/.."
python3 ./simulator.py --config-name gtg_sv/mnist.yaml
