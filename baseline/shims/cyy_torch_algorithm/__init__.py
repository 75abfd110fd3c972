"""Stand-in for the private ``cyy_torch_algorithm`` (quantizers, Shapley engines) in plain
PyTorch / numpy; only what the reference imports at package-import time plus fed_paq / GTG."""
