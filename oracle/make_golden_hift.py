"""TEST INFRASTRUCTURE ONLY -- HiFT / end-to-end / streaming golden vectors (filled in with the HiFT kernels)."""


def run(what, flow):
    pass
