def mark_used(*_a):
    return None
def count_none(*args):
    return sum(1 for a in args if a is None)
