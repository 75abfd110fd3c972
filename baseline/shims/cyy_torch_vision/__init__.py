"""Import-time registration of datasets/models happens in cyy_torch_toolbox here."""
