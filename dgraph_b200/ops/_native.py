"""placeholder; replaced below"""
def available():
    return False
