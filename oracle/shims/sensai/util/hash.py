import hashlib, pickle
def pickle_hash(o, **_k):
    return hashlib.sha1(pickle.dumps(o)).hexdigest()
