"""Minimal labelled arrays: the subset of the xarray interface that vcm.cubedsphere relies on.

xarray is not installed in this image, and the reference's operator API
(/root/reference/external/vcm/vcm/cubedsphere/__init__.py) takes and returns ``xr.DataArray`` /
``xr.Dataset``.  ``fv3net_b200.cubedsphere`` is written against the attributes both libraries
share (``dims, values, coords, attrs, name, sizes`` and the constructor signature
``DataArray(data, coords=, dims=, name=, attrs=)`` / ``Dataset(data_vars, coords=, attrs=)``), so
real xarray objects work where xarray exists and these classes work everywhere.  This is host
metadata only (dimension names, 1-D coordinates, attributes): no arithmetic of the hot path
happens here.
"""
import numpy as np


def _as_coord(name, value, owner_dims=None):
    """-> (dims, ndarray) of one coordinate."""
    if isinstance(value, DataArray):
        return tuple(value.dims), np.asarray(value.values)
    if hasattr(value, "dims") and hasattr(value, "values"):  # a foreign labelled array
        return tuple(value.dims), np.asarray(value.values)
    if isinstance(value, tuple) and len(value) == 2 and isinstance(value[0], (str, tuple, list)):
        dims = (value[0],) if isinstance(value[0], str) else tuple(value[0])
        return dims, np.asarray(value[1])
    arr = np.asarray(value)
    return ((name,) if arr.ndim == 1 else ()), arr


class _Coords(dict):
    """name -> DataArray (1-D or scalar), like ``xr.DataArray.coords``."""

    def to_plain(self):
        return {k: (tuple(v.dims), np.asarray(v.values)) for k, v in self.items()}


class DataArray:
    def __init__(self, data, coords=None, dims=None, name=None, attrs=None):
        self.values = np.asarray(data)
        if dims is None:
            dims = tuple(f"dim_{i}" for i in range(self.values.ndim))
        self.dims = (dims,) if isinstance(dims, str) else tuple(dims)
        if len(self.dims) != self.values.ndim:
            raise ValueError(f"different number of dimensions on data ({self.values.ndim}) and dims ({len(self.dims)})")
        self.name = name
        self.attrs = dict(attrs) if attrs else {}
        self._coords = {}
        if coords is not None:
            if hasattr(coords, "keys"):
                items = [(k, coords[k]) for k in coords.keys()]
            else:  # positional: one per dimension
                items = list(zip(self.dims, coords))
            for k, v in items:
                cd, cv = _as_coord(k, v)
                self._set_coord(k, cd, cv)

    # -- construction helpers ------------------------------------------------------------
    def _set_coord(self, name, cdims, cvalues):
        for d, n in zip(cdims, cvalues.shape):
            if d not in self.dims:
                raise ValueError(f"coordinate {name!r} has dimension {d!r} not on the array")
            if self.sizes[d] != n:
                raise ValueError(f"conflicting sizes for dimension {d!r}: coordinate {name!r} has {n}, array has {self.sizes[d]}")
        self._coords[name] = (tuple(cdims), np.asarray(cvalues))

    def _replace(self, data=None, dims=None, coords=None, name="__keep__", attrs=None):
        out = DataArray.__new__(DataArray)
        out.values = self.values if data is None else np.asarray(data)
        out.dims = self.dims if dims is None else tuple(dims)
        out.name = self.name if name == "__keep__" else name
        out.attrs = dict(self.attrs if attrs is None else attrs)
        out._coords = dict(self._coords if coords is None else coords)
        return out

    # -- xarray-like attributes -------------------------------------------------------------
    @property
    def data(self):
        return self.values

    @property
    def shape(self):
        return self.values.shape

    @property
    def ndim(self):
        return self.values.ndim

    @property
    def dtype(self):
        return self.values.dtype

    @property
    def sizes(self):
        return dict(zip(self.dims, self.values.shape))

    @property
    def chunks(self):
        return None

    @property
    def coords(self):
        return _Coords({k: DataArray(v, dims=d, name=k, coords={k: (d, v)} if d == (k,) else None)
                        for k, (d, v) in self._coords.items()})

    def get_axis_num(self, dim):
        return self.dims.index(dim)

    def __array__(self, dtype=None, copy=None):
        return np.asarray(self.values, dtype=dtype)

    def __len__(self):
        return len(self.values)

    def __repr__(self):
        c = ", ".join(f"{k}{list(d)}" for k, (d, _) in self._coords.items())
        return f"<fv3net_b200.DataArray {self.name!r} {self.sizes} {self.dtype} coords=({c}) attrs={self.attrs}>\n{self.values!r}"

    # -- indexing -----------------------------------------------------------------------------
    def __getitem__(self, key):
        if isinstance(key, str):
            if key in self._coords:
                return self.coords[key]
            if key in self.dims:  # xarray: a dimension without a coordinate reads as arange
                return DataArray(np.arange(self.sizes[key]), dims=(key,), name=key)
            raise KeyError(key)
        if not isinstance(key, tuple):
            key = (key,)
        return self.isel({d: k for d, k in zip(self.dims, key)})

    def __setitem__(self, key, value):
        if not isinstance(key, str):
            raise TypeError("only coordinate assignment by name is supported")
        cd, cv = _as_coord(key, value)
        self._set_coord(key, cd, cv)

    def isel(self, indexers=None, **kw):
        indexers = dict(indexers or {}, **kw)
        idx, dims = [], []
        for d in self.dims:
            k = indexers.get(d, slice(None))
            idx.append(k)
            if not isinstance(k, (int, np.integer)):
                dims.append(d)
        coords = {}
        for name, (cd, cv) in self._coords.items():
            ck = tuple(indexers.get(d, slice(None)) for d in cd)
            nd = tuple(d for d, k in zip(cd, ck) if not isinstance(k, (int, np.integer)))
            coords[name] = (nd, cv[ck])
        return self._replace(self.values[tuple(idx)], dims, coords)

    def transpose(self, *dims):
        if not dims:
            dims = self.dims[::-1]
        if set(dims) != set(self.dims):
            raise ValueError(f"{dims} must be a permutation of {self.dims}")
        return self._replace(self.values.transpose([self.dims.index(d) for d in dims]), dims)

    # -- metadata -----------------------------------------------------------------------------
    def rename(self, new_name_or_name_dict=None, **names):
        if hasattr(new_name_or_name_dict, "keys") or names:
            m = dict(new_name_or_name_dict or {}, **names)
            dims = tuple(m.get(d, d) for d in self.dims)
            coords = {m.get(k, k): (tuple(m.get(d, d) for d in cd), cv) for k, (cd, cv) in self._coords.items()}
            return self._replace(dims=dims, coords=coords)
        return self._replace(name=new_name_or_name_dict)

    def assign_coords(self, coords=None, **kw):
        out = self._replace()
        for k, v in dict(coords or {}, **kw).items():
            out[k] = v
        return out

    def assign_attrs(self, *args, **kw):
        out = self._replace()
        for a in args:
            out.attrs.update(a)
        out.attrs.update(kw)
        return out

    def drop_vars(self, names):
        names = [names] if isinstance(names, str) else list(names)
        return self._replace(coords={k: v for k, v in self._coords.items() if k not in names})

    drop = drop_vars

    def copy(self, deep=True, data=None):
        return self._replace(np.array(self.values) if (deep and data is None) else data)

    def astype(self, dtype):
        return self._replace(self.values.astype(dtype))

    def to_dataset(self, name=None):
        name = name if name is not None else self.name
        if name is None:
            raise ValueError("unable to convert unnamed DataArray to a Dataset without providing an explicit name")
        return Dataset({name: self})

    def where(self, cond, other=np.nan):
        c = cond.values if hasattr(cond, "values") else np.asarray(cond)
        return self._replace(np.where(c, self.values, other))

    # -- arithmetic (broadcast by dimension name: self's dims first, then the other's) ----------
    def _binary(self, other, op, reflexive=False):
        if isinstance(other, Dataset):
            return NotImplemented
        if hasattr(other, "dims") and hasattr(other, "values"):
            dims = self.dims + tuple(d for d in other.dims if d not in self.dims)
            a = _expand(self.values, self.dims, dims)
            b = _expand(np.asarray(other.values), tuple(other.dims), dims)
            coords = dict(self._coords)
            oc = other._coords if isinstance(other, DataArray) else {}
            for k, v in oc.items():
                coords.setdefault(k, v)
            name = self.name if getattr(other, "name", None) == self.name else None
        else:
            dims, a, b, coords, name = self.dims, self.values, other, self._coords, self.name
        with np.errstate(all="ignore"):
            data = op(b, a) if reflexive else op(a, b)
        return DataArray._replace(self, data, dims, coords, name=name, attrs={})

    def __add__(self, o): return self._binary(o, np.add)
    def __radd__(self, o): return self._binary(o, np.add, True)
    def __sub__(self, o): return self._binary(o, np.subtract)
    def __rsub__(self, o): return self._binary(o, np.subtract, True)
    def __mul__(self, o): return self._binary(o, np.multiply)
    def __rmul__(self, o): return self._binary(o, np.multiply, True)
    def __truediv__(self, o): return self._binary(o, np.true_divide)
    def __rtruediv__(self, o): return self._binary(o, np.true_divide, True)
    def __floordiv__(self, o): return self._binary(o, np.floor_divide)
    def __lt__(self, o): return self._binary(o, np.less)
    def __gt__(self, o): return self._binary(o, np.greater)
    def __neg__(self): return self._replace(-self.values)

    # -- comparison (xr.testing.assert_identical semantics + dtype) ------------------------------
    def equals(self, other):
        return (isinstance(other, DataArray) and self.dims == other.dims and self.shape == other.shape
                and _array_equal(self.values, other.values) and _coords_equal(self._coords, other._coords))

    def identical(self, other):
        return self.equals(other) and self.name == other.name and _attrs_equal(self.attrs, other.attrs)


def _expand(values, dims, out_dims):
    order = [dims.index(d) for d in out_dims if d in dims]
    v = values.transpose(order) if order != list(range(len(dims))) else values
    shape = tuple(values.shape[dims.index(d)] if d in dims else 1 for d in out_dims)
    return v.reshape(shape)


def _array_equal(a, b):
    if a.shape != b.shape:
        return False
    if a.dtype.kind in "fc" or b.dtype.kind in "fc":
        return bool(np.array_equal(a, b, equal_nan=True))
    return bool(np.array_equal(a, b))


def _coords_equal(a, b):
    return set(a) == set(b) and all(a[k][0] == b[k][0] and _array_equal(a[k][1], b[k][1]) for k in a)


def _attrs_equal(a, b):
    if set(a) != set(b):
        return False
    for k in a:
        x, y = a[k], b[k]
        if isinstance(x, np.ndarray) or isinstance(y, np.ndarray):
            if not np.array_equal(x, y):
                return False
        elif x != y:
            return False
    return True


class Dataset:
    """name -> DataArray with shared dimension sizes, like ``xr.Dataset``."""

    def __init__(self, data_vars=None, coords=None, attrs=None):
        self._vars = {}
        self._coords = {}
        self.attrs = dict(attrs) if attrs else {}
        if coords is not None:
            for k in coords.keys():
                self._coords[k] = _as_coord(k, coords[k])
        for k, v in (data_vars or {}).items():
            self[k] = v

    def __setitem__(self, name, value):
        if isinstance(value, tuple) and not isinstance(value, DataArray):
            value = DataArray(value[1], dims=value[0], attrs=value[2] if len(value) > 2 else None)
        if not isinstance(value, DataArray):
            if hasattr(value, "dims") and hasattr(value, "values"):
                value = DataArray(value.values, dims=value.dims, attrs=getattr(value, "attrs", None))
            else:
                value = DataArray(value)
        sizes = self.sizes
        for d, n in value.sizes.items():
            if d in sizes and sizes[d] != n:
                raise ValueError(f"conflicting sizes for dimension {d!r}: {n} vs {sizes[d]}")
        for k, cv in value._coords.items():
            self._coords.setdefault(k, cv)
        self._vars[name] = value._replace(name=name, coords={})

    def _var_coords(self, v):
        return {k: cv for k, cv in self._coords.items() if all(d in v.dims for d in cv[0])}

    def __getitem__(self, name):
        if name in self._vars:
            v = self._vars[name]
            return v._replace(coords=self._var_coords(v))
        if name in self._coords:
            d, v = self._coords[name]
            return DataArray(v, dims=d, name=name, coords={name: (d, v)} if d == (name,) else None)
        if name in self.sizes:
            return DataArray(np.arange(self.sizes[name]), dims=(name,), name=name)
        raise KeyError(name)

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name)

    def __iter__(self):
        return iter(self._vars)

    def __contains__(self, name):
        return name in self._vars or name in self._coords

    def __len__(self):
        return len(self._vars)

    def keys(self):
        return self._vars.keys()

    def items(self):
        return [(k, self[k]) for k in self._vars]

    @property
    def data_vars(self):
        return {k: self[k] for k in self._vars}

    @property
    def variables(self):
        return list(self._vars) + list(self._coords)

    @property
    def coords(self):
        return _Coords({k: self[k] for k in self._coords})

    @property
    def sizes(self):
        out = {}
        for v in self._vars.values():
            out.update(v.sizes)
        for cd, cv in self._coords.values():
            out.update(dict(zip(cd, cv.shape)))
        return out

    dims = sizes

    def __repr__(self):
        body = "\n".join(f"  {k}: {v.sizes} {v.dtype}" for k, v in self._vars.items())
        return f"<fv3net_b200.Dataset sizes={self.sizes} coords={list(self._coords)} attrs={self.attrs}>\n{body}"

    def _rebuild(self, variables, coords=None, attrs=None):
        out = Dataset(attrs=self.attrs if attrs is None else attrs)
        out._coords = dict(self._coords if coords is None else coords)
        for k, v in variables.items():
            out[k] = v
        return out

    def map(self, func, keep_attrs=False, args=(), **kwargs):
        out = Dataset(attrs=self.attrs if keep_attrs else None)
        for k in self._vars:
            out[k] = func(self[k], *args, **kwargs)
        return out

    apply = map

    def copy(self, deep=True):
        return self._rebuild({k: self[k].copy(deep) for k in self._vars})

    def assign_coords(self, coords=None, **kw):
        out = self._rebuild({k: v for k, v in self._vars.items()})
        for k, v in dict(coords or {}, **kw).items():
            out._coords[k] = _as_coord(k, v)
        return out

    def assign_attrs(self, *args, **kw):
        out = self._rebuild({k: v for k, v in self._vars.items()})
        for a in args:
            out.attrs.update(a)
        out.attrs.update(kw)
        return out

    def assign(self, variables=None, **kw):
        out = self._rebuild({k: v for k, v in self._vars.items()})
        for k, v in dict(variables or {}, **kw).items():
            out[k] = v
        return out

    def rename(self, name_dict=None, **names):
        m = dict(name_dict or {}, **names)
        coords = {m.get(k, k): (tuple(m.get(d, d) for d in cd), cv) for k, (cd, cv) in self._coords.items()}
        out = Dataset(attrs=self.attrs)
        out._coords = coords
        for k, v in self._vars.items():
            out[m.get(k, k)] = v.rename({d: m[d] for d in v.dims if d in m})
        return out

    def isel(self, indexers=None, **kw):
        indexers = dict(indexers or {}, **kw)
        out = Dataset(attrs=self.attrs)
        for k in self._vars:
            v = self[k]
            out[k] = v.isel({d: i for d, i in indexers.items() if d in v.dims})
        for k, (cd, cv) in self._coords.items():
            if k not in out._coords:
                ck = tuple(indexers.get(d, slice(None)) for d in cd)
                nd = tuple(d for d, kk in zip(cd, ck) if not isinstance(kk, (int, np.integer)))
                out._coords[k] = (nd, cv[ck])
        return out

    def drop_vars(self, names):
        names = [names] if isinstance(names, str) else list(names)
        out = Dataset(attrs=self.attrs)
        out._coords = {k: v for k, v in self._coords.items() if k not in names}
        for k, v in self._vars.items():
            if k not in names:
                out._vars[k] = v
        return out

    drop = drop_vars

    def equals(self, other):
        return (isinstance(other, Dataset) and set(self._vars) == set(other._vars)
                and all(self._vars[k].equals(other._vars[k]) for k in self._vars)
                and _coords_equal(self._coords, other._coords))

    def identical(self, other):
        return (self.equals(other) and _attrs_equal(self.attrs, other.attrs)
                and all(_attrs_equal(self._vars[k].attrs, other._vars[k].attrs) for k in self._vars))


def merge(objects):
    out = Dataset()
    for o in objects:
        if isinstance(o, DataArray):
            o = o.to_dataset()
        for k in o:
            out[k] = o[k]
        for k, v in o._coords.items():
            out._coords.setdefault(k, v)
    return out


def is_dataset(obj):
    return hasattr(obj, "data_vars")


def assert_identical_including_dtype(a, b):
    """external/vcm/vcm/xarray_utils.py:85-100 for these classes (and for xarray objects)."""
    assert a.identical(b), f"not identical:\n{a!r}\n{b!r}"
    if is_dataset(a):
        for k in a.variables:
            assert a[k].dtype == b[k].dtype, f"dtype of {k!r}: {a[k].dtype} != {b[k].dtype}"
    else:
        assert a.dtype == b.dtype, f"dtype: {a.dtype} != {b.dtype}"
        for k in a.coords:
            assert a.coords[k].dtype == b.coords[k].dtype, f"dtype of coordinate {k!r}"
