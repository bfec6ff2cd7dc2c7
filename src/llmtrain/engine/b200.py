"""placeholder until the native engine lands"""


def supports(model):  # noqa: ANN001, ANN201
    return False
