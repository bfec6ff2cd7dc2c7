"""placeholder"""


def shutdown() -> None:
    return None
