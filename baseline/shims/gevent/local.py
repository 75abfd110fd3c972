import threading


class local(threading.local):
    pass
