class Timer:
    def __init__(self, *a, **k):
        self.last = 0.0

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
