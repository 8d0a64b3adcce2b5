"""Model classes with the attribute names check-gpu-node.py reads (:173-177, :183-187, :203-211)."""


class _Model:
    _fields = ()

    def __init__(self, **kw):
        for f in self._fields:
            setattr(self, f, kw.pop(f, None))
        if kw:
            raise TypeError(f"{type(self).__name__}: unexpected fields {sorted(kw)}")

    def __repr__(self):
        return f"{type(self).__name__}({', '.join(f'{f}={getattr(self, f)!r}' for f in self._fields)})"


class V1ObjectMeta(_Model):
    _fields = ("name", "labels", "annotations")


class V1Taint(_Model):
    _fields = ("key", "value", "effect", "time_added")


class V1NodeSpec(_Model):
    _fields = ("taints", "unschedulable", "provider_id")


class V1NodeCondition(_Model):
    _fields = ("type", "status", "reason", "message", "last_heartbeat_time", "last_transition_time")


class V1NodeStatus(_Model):
    _fields = ("capacity", "allocatable", "conditions", "addresses", "node_info")


class V1Node(_Model):
    _fields = ("metadata", "spec", "status", "api_version", "kind")


class V1NodeList(_Model):
    _fields = ("items", "metadata", "api_version", "kind")


class CoreV1Api:
    """`list_node()` returns whatever fleet the test installed with `set_fleet` (class-wide)."""

    _fleet = None

    def __init__(self, api_client=None):
        self.api_client = api_client

    @classmethod
    def set_fleet(cls, items):
        cls._fleet = items

    def list_node(self, **kwargs):
        return V1NodeList(items=type(self)._fleet)
