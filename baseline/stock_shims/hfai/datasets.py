def ImageNet(*a, **k):
    raise RuntimeError("ffrecord ImageNet is not available; the bench feeds synthetic batches")
