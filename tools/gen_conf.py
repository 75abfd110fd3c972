#!/usr/bin/env python3
"""Regenerates ``distributed_learning_simulator_b200/conf/**`` from one compact table.

The experiment matrix (which algorithm runs which model on which dataset with how many
workers/rounds) is the reference's public interface -- its scripts address these files
by name (/root/reference/test.sh:2-10, fed_obd_train.sh:1-3) -- so file names, keys and
values are kept; the files themselves are emitted by this generator.
"""
import os

import yaml

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "distributed_learning_simulator_b200", "conf")

IMDB_MODEL = {"max_len": 300, "word_vector_name": "glove.6B.100d", "num_encoder_layer": 2, "d_model": 100, "nhead": 5}
DATA = {
    "mnist": dict(dataset_name="MNIST", model_name="LeNet5", learning_rate=0.01, round=20, epoch=2),
    "cifar10": dict(dataset_name="CIFAR10", model_name="densenet40", learning_rate=0.1, round=100, epoch=5),
    "cifar100": dict(dataset_name="CIFAR100", model_name="densenet40", learning_rate=0.1, round=100, epoch=5),
    "imdb": dict(dataset_name="imdb", model_name="TransformerClassificationModel", learning_rate=0.01, round=100,
                 epoch=5, dataset_kwargs={"max_len": 300}, model_kwargs=IMDB_MODEL),
    "imagenet": dict(dataset_name="ImageNet", model_name="Resnet50", learning_rate=0.01, round=100, epoch=5,
                     batch_size=128),
}
SGD = dict(optimizer_name="SGD", learning_rate_scheduler_name="CosineAnnealingLR", batch_size=64, worker_number=10)


def cv(algo, data, **extra):
    conf = dict(SGD)
    conf.update(DATA[data])
    conf["distributed_algorithm"] = algo
    conf.update(extra)
    return conf


def obd(data, workers, chosen, weight, dropout, algo="fed_obd"):
    extra = dict(worker_number=workers,
                 algorithm_kwargs={"second_phase_epoch": 10, "dropout_rate": dropout, "random_client_number": chosen})
    if weight is not None:
        extra["endpoint_kwargs"] = {"server": {"weight": weight}, "worker": {"weight": weight}}
    return cv(algo, data, **extra)


def graph(algo, dataset, model, workers, rounds, epoch=1, lr=0.001, wd=0, **extra):
    conf = dict(dataset_name=dataset, model_name=model, distributed_algorithm=algo, epoch=epoch,
                learning_rate=lr, weight_decay=wd, round=rounds, worker_number=workers)
    conf.update(extra)
    return conf


def drop_kw(chosen):
    return {"algorithm_kwargs": {"dropout_rate": 0.3, "random_client_number": chosen}}


FILES = {"global.yaml": dict(cache_transforms="cpu", log_level="INFO", save_performance_metric=False,
                             use_slow_performance_metrics=True, merge_validation_to_training_set=False,
                             use_amp=False)}
for d in ("mnist", "cifar10", "cifar100", "imagenet"):
    FILES[f"fed_avg/{d}.yaml"] = cv("fed_avg", d)
FILES["fed_avg/imdb.yaml"] = cv("fed_avg", "imdb", dataset_kwargs={"max_len": 300, "tokenizer": {"type": "spacy"}})
for d in ("cifar10", "cifar100", "imdb"):
    FILES[f"fed_dropout_avg/{d}.yaml"] = cv("fed_dropout_avg", d, **drop_kw(5))
    FILES[f"fed_paq/{d}.yaml"] = cv("fed_paq", d, algorithm_kwargs={"random_client_number": 5})
    FILES[f"smafd/{d}.yaml"] = cv("single_model_afd", d, **drop_kw(5))
    FILES[f"sign_sgd/{d}.yaml"] = cv("sign_SGD", d, distribute_init_parameters=False, round=1, epoch=100)
    FILES[f"fed_obd/{d}.yaml"] = obd(d, 10, 5, 0.001 if d == "imdb" else 0.01, 0.9)
    big = 50 if d == "cifar100" else 100
    FILES[f"large_scale/fed_avg/{d}.yaml"] = cv("fed_avg", d, worker_number=big)
    FILES[f"large_scale/fed_dropout_avg/{d}.yaml"] = cv("fed_dropout_avg", d, worker_number=big, **drop_kw(big // 2))
    FILES[f"large_scale/fed_paq/{d}.yaml"] = cv("fed_paq", d, worker_number=big,
                                                algorithm_kwargs={"random_client_number": big // 2})
    FILES[f"large_scale/smafd/{d}.yaml"] = cv("single_model_afd", d, worker_number=big, **drop_kw(big // 2))
    FILES[f"large_scale/fed_obd/{d}.yaml"] = obd(d, big, big // 2, 0.0001 if d == "imdb" else 0.001, 0.3)
del FILES["sign_sgd/cifar100.yaml"]["batch_size"]  # the reference file omits it
FILES["fed_obd_sq/cifar100.yaml"] = obd("cifar100", 10, 5, None, 0.9, algo="fed_obd_sq")
FILES["large_scale/fed_obd/cifar100_sq.yaml"] = obd("cifar100", 50, 25, None, 0.3, algo="fed_obd_sq")
FILES["gtg_sv/mnist.yaml"] = cv("GTG_shapley_value", "mnist")
FILES["gtg_sv/cifar10.yaml"] = cv("GTG_shapley_value", "cifar10")
FILES["gtg_sv/cifar100.yaml"] = cv("GTG_shapley_value", "cifar100")
FILES["gtg_sv/imdb.yaml"] = cv("fed_avg", "imdb")  # sic: the reference file names fed_avg
FILES["multiround_sv/cifar10.yaml"] = cv("multiround_shapley_value", "cifar10")
FILES["multiround_sv/cifar100.yaml"] = cv("GTG_shapley_value", "cifar100")  # sic
FILES["hierarchical_sv/cifar10.yaml"] = cv("GTG_shapley_value", "cifar10")  # sic
FILES["hierarchical_sv/mnist.yaml"] = cv("Hierarchical_shapley_value", "mnist", worker_number=6, round=2, epoch=1,
                                         algorithm_kwargs={"part_number": 3, "vp_size": 3})
NN = {"extra_hyper_parameters": {"num_neighbor": 10}}
FILES["fed_gnn/cs.yaml"] = graph("fed_gnn", "Coauthor_CS", "TwoGCN", 50, 50, use_amp=False,
                                 algorithm_kwargs={"share_feature": True, "batch_number": 10, "edge_drop_rate": 0.99,
                                                   "num_neighbor": 10})
FILES["fed_gnn/yelp.yaml"] = graph("fed_gcn", "Yelp", "TwoGCN", 50, 50, use_amp=False, algorithm_kwargs={
    "share_feature": False, "batch_number": 10, "edge_drop_rate": 0.99}, **NN)
FILES["fed_gnn/amazonproduct.yaml"] = graph("fed_gcn", "AmazonProducts", "TwoGCN", 50, 50, use_amp=False,
                                            algorithm_kwargs={"share_feature": False, "batch_number": 20,
                                                              "edge_drop_rate": 0.99}, **NN)
FILES["fed_aas/PubMed.yaml"] = graph("fed_aas", "PubMed", "SimpleGCN", 100, 1000, epoch=5)
FILES["fed_aas/cora.yaml"] = graph("fed_aas", "Cora", "SimpleGCN", 10, 10, epoch=10, lr=0.01, wd=0.01, use_amp=True,
                                   parallel_number=2)
FILES["fed_aas/dblp.yaml"] = graph("fed_aas", "CitationFull", "OneGCN", 100, 50, use_amp=False,
                                   dataset_kwargs={"name": "DBLP"},
                                   algorithm_kwargs={"share_feature": False, "batch_number": 1})
for name, ds, rounds in (("reddit", "Reddit", 100), ("yelp", "Yelp", 50)):
    FILES[f"fed_aas/{name}.yaml"] = graph("fed_aas", ds, "TwoGCN", 50, rounds, use_amp=False, algorithm_kwargs={
        "share_feature": False, "batch_number": 10, "edge_drop_rate": 0.99}, **NN)

ORDER = ["dataset_name", "dataset_kwargs", "model_name", "model_kwargs", "distributed_algorithm", "worker_number",
         "round", "epoch", "batch_size", "optimizer_name", "learning_rate", "learning_rate_scheduler_name"]


def main() -> None:
    for rel, conf in sorted(FILES.items()):
        path = os.path.join(ROOT, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        keys = [k for k in ORDER if k in conf] + sorted(k for k in conf if k not in ORDER)
        with open(path, "wt", encoding="utf8") as f:
            f.write(f"# {rel} -- generated by tools/gen_conf.py; edit the table there\n")
            yaml.safe_dump({k: conf[k] for k in keys}, f, sort_keys=False, default_flow_style=False)
    print(f"wrote {len(FILES)} files under {os.path.normpath(ROOT)}")


if __name__ == "__main__":
    main()
