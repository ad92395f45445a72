def breakpoint(*_a, **_k):
    pass
