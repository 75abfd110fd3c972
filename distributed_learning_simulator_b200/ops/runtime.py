def available() -> bool:
    return False
