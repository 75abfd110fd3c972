#!/bin/sh
set -e
cd "$(dirname "$0")/.."
SMALL="++%s.dataset_kwargs.train_size=2000 ++%s.dataset_kwargs.test_size=400"
# Fed dropout avg
python3 ./simulator.py --config-name fed_dropout_avg/cifar100.yaml ++fed_dropout_avg.round=1 ++fed_dropout_avg.epoch=1 ++fed_dropout_avg.worker_number=2 ++fed_dropout_avg.debug=True ++fed_dropout_avg.algorithm_kwargs.random_client_number=2 $(printf "$SMALL" fed_dropout_avg fed_dropout_avg)
python3 ./simulator.py --config-name fed_paq/cifar100.yaml ++fed_paq.round=1 ++fed_paq.epoch=1 ++fed_paq.worker_number=2 ++fed_paq.debug=True ++fed_paq.algorithm_kwargs.random_client_number=2 $(printf "$SMALL" fed_paq fed_paq)
# the four methods the reference names in conf/ but does not implement
python3 ./simulator.py --config-name sign_sgd/cifar10.yaml ++sign_sgd.epoch=1 ++sign_sgd.worker_number=2 $(printf "$SMALL" sign_sgd sign_sgd)
python3 ./simulator.py --config-name smafd/cifar10.yaml ++smafd.round=1 ++smafd.epoch=1 ++smafd.worker_number=4 ++smafd.algorithm_kwargs.random_client_number=2 $(printf "$SMALL" smafd smafd)
python3 ./simulator.py --config-name fed_aas/cora.yaml ++fed_aas.round=1 ++fed_aas.epoch=1 ++fed_aas.worker_number=2
python3 ./simulator.py --config-name hierarchical_sv/mnist.yaml ++hierarchical_sv.round=1 $(printf "$SMALL" hierarchical_sv hierarchical_sv)
