"""placeholder (filled in below)"""
class FusedMPMLP:  # noqa: D101
    @classmethod
    def from_sequential(cls, model, **kw):
        raise NotImplementedError
