class GraphDatasetUtil:  # only the name is needed to import the reference's graph worker
    pass
