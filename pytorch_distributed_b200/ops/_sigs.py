def declare(L):
    pass
