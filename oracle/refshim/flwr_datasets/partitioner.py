class IidPartitioner:
    def __init__(self, *a, **kw):
        pass


class DirichletPartitioner(IidPartitioner):
    pass
