"""Minimal stand-in for the slice of pyspark.ml that spark.py touches, used ONLY when pyspark cannot be imported
(this image has no pyspark and no JVM).  It mirrors pyspark 3.1.2 semantics for: Param / Params (uid, defaults, set,
getOrDefault, copy(extra) keeping the uid, extractParamMap, explainParams), TypeConverters.toInt (integral floats
accepted, `ML 08 - Hyperopt.py:97`), Estimator.fit(dataset, params) / fitMultiple, Model.transform(dataset, params),
and a LocalFrame that carries columns plus the `ml_attr` metadata VectorAssembler attaches to the features column.
"""
from __future__ import annotations

import copy as _copy
import threading
import uuid

import numpy as np


class TypeConverters:
    @staticmethod
    def toInt(v):
        if isinstance(v, bool):
            raise TypeError(f"Could not convert {v!r} to int")
        if isinstance(v, (int, np.integer)):
            return int(v)
        if isinstance(v, (float, np.floating)) and float(v).is_integer():
            return int(v)
        raise TypeError(f"Could not convert {v!r} to int")

    @staticmethod
    def toFloat(v):
        if isinstance(v, bool) or not isinstance(v, (int, float, np.integer, np.floating)):
            raise TypeError(f"Could not convert {v!r} to float")
        return float(v)

    @staticmethod
    def toListFloat(v):
        if v is None:
            return None
        return [TypeConverters.toFloat(x) for x in list(v)]

    @staticmethod
    def toString(v):
        if isinstance(v, str):
            return v
        raise TypeError(f"Could not convert {v!r} to string type")

    @staticmethod
    def toBoolean(v):
        if isinstance(v, bool):
            return v
        raise TypeError(f"Boolean Param requires value of type bool. Found {type(v)}.")

    @staticmethod
    def identity(v):
        return v


class Param:
    def __init__(self, parent, name, doc, typeConverter=None):
        self.parent = parent.uid if hasattr(parent, "uid") else parent
        self.name, self.doc = name, doc
        self.typeConverter = typeConverter or TypeConverters.identity

    def _copy_new_parent(self, parent):
        p = _copy.copy(self)
        p.parent = parent.uid
        return p

    def __hash__(self):
        return hash((self.parent, self.name))

    def __eq__(self, o):
        return isinstance(o, Param) and self.parent == o.parent and self.name == o.name

    def __repr__(self):
        return f"Param(parent={self.parent!r}, name={self.name!r}, doc={self.doc!r})"


class Params:
    def __init__(self):
        self.uid = f"{type(self).__name__}_{uuid.uuid4().hex[:12]}"
        self._paramMap, self._defaultParamMap = {}, {}
        for name in dir(type(self)):
            v = getattr(type(self), name, None)
            if isinstance(v, Param):
                setattr(self, name, v._copy_new_parent(self))

    @property
    def params(self):
        return sorted((getattr(self, n) for n in dir(self) if n != "params" and isinstance(getattr(type(self), n, None), Param)),
                      key=lambda p: p.name)

    def getParam(self, name):
        p = getattr(self, name, None)
        if not isinstance(p, Param):
            raise ValueError(f"Cannot find param with name {name}.")
        return p

    def hasParam(self, name):
        return isinstance(getattr(self, name, None), Param)

    def _resolve(self, param):
        if isinstance(param, Param):
            if param.parent != self.uid:
                raise ValueError(f"Param {param} does not belong to {self.uid}.")
            return getattr(self, param.name)
        return self.getParam(param)

    def isSet(self, param):
        return self._resolve(param) in self._paramMap

    def isDefined(self, param):
        p = self._resolve(param)
        return p in self._paramMap or p in self._defaultParamMap

    def getOrDefault(self, param):
        p = self._resolve(param)
        if p in self._paramMap:
            return self._paramMap[p]
        return self._defaultParamMap[p]

    def _set(self, **kw):
        for k, v in kw.items():
            p = self.getParam(k)
            if v is not None:
                try:
                    v = p.typeConverter(v)
                except TypeError as e:
                    raise TypeError(f'Invalid param value given for param "{p.name}". {e}')
            self._paramMap[p] = v
        return self

    def set(self, param, value):
        return self._set(**{self._resolve(param).name: value})

    def _setDefault(self, **kw):
        for k, v in kw.items():
            p = self.getParam(k)
            self._defaultParamMap[p] = v if v is None else p.typeConverter(v)
        return self

    def extractParamMap(self, extra=None):
        m = dict(self._defaultParamMap)
        m.update(self._paramMap)
        if extra:
            m.update(extra)
        return m

    def explainParam(self, param):
        p = self._resolve(param)
        vals = []
        if p in self._defaultParamMap:
            vals.append(f"default: {self._defaultParamMap[p]}")
        if p in self._paramMap:
            vals.append(f"current: {self._paramMap[p]}")
        return f"{p.name}: {p.doc} ({', '.join(vals) if vals else 'undefined'})"

    def explainParams(self):
        return "\n".join(self.explainParam(p) for p in self.params)

    def copy(self, extra=None):
        that = _copy.copy(self)
        that._paramMap = dict(self._paramMap)
        that._defaultParamMap = dict(self._defaultParamMap)
        return self._copyValues(that, extra)

    def _copyValues(self, to, extra=None):
        pm = dict(self._paramMap)
        if extra:
            pm.update(extra)
        for p, v in pm.items():
            if isinstance(p, Param) and p.parent == self.uid and to.hasParam(p.name):     # routed by owning uid
                to._set(**{p.name: v})
        return to


class Estimator(Params):
    def fit(self, dataset, params=None):
        if params is None:
            params = {}
        if isinstance(params, (list, tuple)):
            models = [None] * len(params)
            for i, m in self.fitMultiple(dataset, params):
                models[i] = m
            return models
        if isinstance(params, dict):
            return self.copy(params)._fit(dataset) if params else self._fit(dataset)
        raise TypeError("Params must be either a param map or a list/tuple of param maps")

    def fitMultiple(self, dataset, paramMaps):
        est = self.copy()
        lock, it = threading.Lock(), iter(range(len(paramMaps)))

        class _It:
            def __iter__(s):
                return s

            def __next__(s):
                with lock:
                    i = next(it)
                return i, est.fit(dataset, paramMaps[i])
            next = __next__
        return _It()


class Transformer(Params):
    def transform(self, dataset, params=None):
        if params:
            return self.copy(params)._transform(dataset)
        return self._transform(dataset)


class Model(Transformer):
    pass


class SparseVector:
    def __init__(self, size, indices, values):
        self.size, self.indices, self.values = int(size), np.asarray(indices, np.int32), np.asarray(values, np.float64)

    def toArray(self):
        a = np.zeros(self.size)
        a[self.indices] = self.values
        return a

    def __iter__(self):
        return iter(self.toArray())

    def __len__(self):
        return self.size

    def __getitem__(self, i):
        return self.toArray()[i]

    def __repr__(self):
        return f"SparseVector({self.size}, {{{', '.join(f'{i}: {v}' for i, v in zip(self.indices, self.values))}}})"


class DenseVector:
    def __init__(self, values):
        self.values = np.asarray(values, np.float64)

    def toArray(self):
        return self.values

    def __iter__(self):
        return iter(self.values)

    def __len__(self):
        return len(self.values)

    def __getitem__(self, i):
        return self.values[i]

    def __eq__(self, other):
        return np.array_equal(self.values, np.asarray(other.toArray() if hasattr(other, "toArray") else other, np.float64))

    def __repr__(self):
        return f"DenseVector([{', '.join(repr(float(v)) for v in self.values)}])"


class _Field:
    def __init__(self, name, metadata):
        self.name, self.metadata = name, metadata


class _Schema:
    def __init__(self, fields):
        self._f = fields

    def __getitem__(self, name):
        return self._f[name]


class LocalFrame:
    """In-memory DataFrame look-alike: named columns (numpy arrays; the features column is a 2-D array) plus per-column
    metadata in the shape Spark uses (`schema[featuresCol].metadata["ml_attr"]`)."""

    def __init__(self, columns: dict, metadata: dict = None, num_partitions: int = 1):
        self._cols = dict(columns)
        self._meta = dict(metadata or {})
        self.num_partitions = num_partitions

    @property
    def columns(self):
        return list(self._cols)

    @property
    def schema(self):
        return _Schema({k: _Field(k, self._meta.get(k, {})) for k in self._cols})

    def count(self):
        return len(next(iter(self._cols.values())))

    def col(self, name):
        return self._cols[name]

    def withColumn(self, name, values, metadata=None):
        c, m = dict(self._cols), dict(self._meta)
        c[name] = values
        if metadata is not None:
            m[name] = metadata
        return LocalFrame(c, m, self.num_partitions)

    def filter_rows(self, mask):
        return LocalFrame({k: v[mask] for k, v in self._cols.items()}, self._meta, self.num_partitions)


def ml_attr_metadata(arity, names=None):
    """The `ml_attr` metadata VectorAssembler writes for a features column with the given per-slot arities (0 = numeric)."""
    attrs = {"numeric": [], "nominal": []}
    for i, a in enumerate(arity):
        nm = names[i] if names else f"f{i}"
        if a and a > 0:
            attrs["nominal"].append({"idx": i, "name": nm, "vals": [str(v) for v in range(int(a))]})
        else:
            attrs["numeric"].append({"idx": i, "name": nm})
    return {"ml_attr": {"attrs": {k: v for k, v in attrs.items() if v}, "num_attrs": len(arity)}}
