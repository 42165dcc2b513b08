def declare(lib):
    pass
