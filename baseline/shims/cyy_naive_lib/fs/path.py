import os


def find_directories(root: str, name: str) -> list:
    return sorted(os.path.join(d, name) for d, subdirs, _ in os.walk(root) if name in subdirs)
