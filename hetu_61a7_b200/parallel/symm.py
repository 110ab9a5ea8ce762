class SymmEngine(object):
    @staticmethod
    def get(comm):
        return None
class FusedDPUpdate(object):
    pass
