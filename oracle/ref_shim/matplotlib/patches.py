class FancyArrowPatch(object):
    pass
