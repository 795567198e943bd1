"""In-process stand-in for the `redis` client package, just wide enough for the reference's RAD traversal
(rad/traverser.py, coordination_service.py, priority_queue.py, visited.py, scored.py, redis_server.py): this container
has neither the `redis` wheel nor a `redis-server` binary.  Test infrastructure only -- installed as
``sys.modules["redis"]`` by tests/test_rad_traversal.py before the reference modules are imported.

One shared keyspace per (host, port), guarded by one lock, values stored as bytes like a real server with
``decode_responses=False``.  The three Lua scripts RAD registers are recognised by the commands they contain."""
import threading


class ConnectionError(Exception):  # noqa: A001 -- mirrors redis.ConnectionError
    pass


class _Exceptions:
    ConnectionError = ConnectionError


exceptions = _Exceptions()
_SERVERS = {}
_LOCK = threading.RLock()


def _b(v):
    if isinstance(v, bytes):
        return v
    if isinstance(v, float):
        return repr(v).encode()
    return str(v).encode()


class _Script:
    def __init__(self, client, text):
        self.client, self.text = client, text

    def __call__(self, keys=(), args=()):
        db = self.client._db
        with _LOCK:
            if "ZRANGE" in self.text:  # priority_queue.py:22-27: pop the lowest score
                z = db.setdefault(_b(keys[0]), {})
                if not z:
                    return None
                member = min(z, key=lambda m: (z[m], m))
                score = z.pop(member)
                return [member, _b(repr(float(score)))]
            if "SISMEMBER" in self.text:  # visited.py:17-24: check-and-insert
                s = db.setdefault(_b(keys[0]), set())
                m = _b(args[0])
                if m in s:
                    return 1
                s.add(m)
                return 0
            if "HEXISTS" in self.text:  # scored.py:37-47: insert once, remember insertion order
                lst = db.setdefault(_b(keys[0]), [])
                scores = db.setdefault(_b(keys[1]), {})
                smiles = db.setdefault(_b(keys[2]), {})
                node = _b(args[0])
                if node not in scores:
                    scores[node] = _b(args[1])
                    smiles[node] = _b(args[2])
                    lst.append(node)
                return None
        raise NotImplementedError(self.text)


class StrictRedis:
    def __init__(self, host="localhost", port=6379, password=None, decode_responses=False, **_):
        with _LOCK:
            self._db = _SERVERS.setdefault((host, int(port)), {})

    # -- connection
    def ping(self):
        return True

    def shutdown(self, save=False):  # noqa: ARG002
        return None

    def close(self):
        return None

    def register_script(self, text):
        return _Script(self, text)

    # -- hashes
    def hset(self, name, key=None, value=None, mapping=None):
        with _LOCK:
            h = self._db.setdefault(_b(name), {})
            n = 0
            items = dict(mapping or {})
            if key is not None:
                items[key] = value
            for k, v in items.items():
                n += _b(k) not in h
                h[_b(k)] = _b(v)
            return n

    def hget(self, name, key):
        with _LOCK:
            return self._db.get(_b(name), {}).get(_b(key))

    def hdel(self, name, *keys):
        with _LOCK:
            h = self._db.get(_b(name), {})
            return sum(h.pop(_b(k), None) is not None for k in keys)

    def hkeys(self, name):
        with _LOCK:
            return list(self._db.get(_b(name), {}).keys())

    def hgetall(self, name):
        with _LOCK:
            return dict(self._db.get(_b(name), {}))

    def hlen(self, name):
        with _LOCK:
            return len(self._db.get(_b(name), {}))

    def hexists(self, name, key):
        with _LOCK:
            return _b(key) in self._db.get(_b(name), {})

    # -- sorted sets, lists, sets, keys
    def zadd(self, name, mapping):
        with _LOCK:
            z = self._db.setdefault(_b(name), {})
            n = 0
            for m, s in mapping.items():
                n += _b(m) not in z
                z[_b(m)] = float(s)
            return n

    def zcard(self, name):
        with _LOCK:
            return len(self._db.get(_b(name), {}))

    def lrange(self, name, start, end):
        with _LOCK:
            lst = self._db.get(_b(name), [])
            return list(lst[start:] if end == -1 else lst[start:end + 1])

    def llen(self, name):
        with _LOCK:
            return len(self._db.get(_b(name), []))

    def scard(self, name):
        with _LOCK:
            return len(self._db.get(_b(name), set()))

    def delete(self, *names):
        with _LOCK:
            return sum(self._db.pop(_b(n), None) is not None for n in names)

    def exists(self, *names):
        with _LOCK:
            return sum(_b(n) in self._db for n in names)

    def keys(self, pattern="*"):
        import fnmatch
        with _LOCK:
            return [k for k in self._db if fnmatch.fnmatch(k.decode(), pattern if isinstance(pattern, str)
                                                           else pattern.decode())]

    def flushall(self):
        with _LOCK:
            self._db.clear()


Redis = StrictRedis
