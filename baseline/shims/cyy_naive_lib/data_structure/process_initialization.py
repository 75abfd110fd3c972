_process_data: dict = {}


def default_initializer(init_arg: dict) -> None:
    _process_data.update(init_arg.get("fun_kwargs", {}))


def get_process_data() -> dict:
    return _process_data
