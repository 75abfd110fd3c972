#!/bin/sh
# the FedOBD paper workloads (reference: fed_obd_train.sh)
cd "$(dirname "$0")/.."
python3 ./simulator.py --config-name large_scale/fed_obd/cifar10.yaml
python3 ./simulator.py --config-name large_scale/fed_obd/cifar100.yaml
python3 ./simulator.py --config-name large_scale/fed_obd/imdb.yaml
