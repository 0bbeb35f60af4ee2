"""Parameter bridge: Spark ML Params <-> backend parameters.

Mirrors the semantics of the reference's ``_CumlClass`` / ``_CumlParams``
(python/src/spark_rapids_ml/params.py:131-555): a Spark-named Param set, a backend-named dict
(``cuml_params``), a name mapping between the two (``""`` = silently ignored, ``None`` = unsupported),
value mappers, conflict detection ("set one or the other"), and the ``num_workers`` /
``float32_inputs`` pseudo-params.  The image has no pyspark, so a small stand-in for
``pyspark.ml.param.Params`` is defined here with the same method names and behaviour that the
estimator code relies on (hasParam/getParam/isSet/isDefined/hasDefault/getOrDefault/_set/
_setDefault/clear/copy/extractParamMap).
"""
from __future__ import annotations

import copy as _copy
import functools
import uuid
from typing import Any, Callable, Dict, List, Optional, Tuple, Union


class Param:
    """Stand-in for pyspark.ml.param.Param: identity is (parent uid, name)."""

    def __init__(self, parent: Any, name: str, doc: str = "", typeConverter: Optional[Callable] = None):
        self.parent = parent.uid if hasattr(parent, "uid") else parent
        self.name = str(name)
        self.doc = doc
        self.typeConverter = typeConverter or (lambda x: x)

    def __repr__(self) -> str:
        return f"{self.parent}__{self.name}"

    def __hash__(self) -> int:
        return hash(str(self))

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, Param) and self.parent == other.parent and self.name == other.name


class TypeConverters:
    @staticmethod
    def toInt(v: Any) -> int:
        if isinstance(v, bool) or not isinstance(v, (int, float)) or int(v) != v:
            raise TypeError(f"Could not convert {v!r} to int")
        return int(v)

    @staticmethod
    def toFloat(v: Any) -> float:
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            raise TypeError(f"Could not convert {type(v)} to float")
        return float(v)

    @staticmethod
    def toString(v: Any) -> str:
        if not isinstance(v, str):
            raise TypeError(f"Could not convert {type(v)} to string type")
        return v

    @staticmethod
    def toListString(v: Any) -> List[str]:
        if isinstance(v, (list, tuple)) and all(isinstance(x, str) for x in v):
            return list(v)
        raise TypeError(f"Could not convert {v!r} to list of strings")


def keyword_only(func: Callable) -> Callable:
    """Same contract as pyspark's decorator: keyword arguments only, saved in ``self._input_kwargs``."""

    @functools.wraps(func)
    def wrapper(self: Any, *args: Any, **kwargs: Any) -> Any:
        if args:
            raise TypeError(f"Method {func.__name__} forces keyword arguments.")
        self._input_kwargs = kwargs
        return func(self, **kwargs)

    return wrapper


class Params:
    """Stand-in for pyspark.ml.param.Params (the subset the estimator / model code uses)."""

    def __init__(self) -> None:
        self.uid = f"{type(self).__name__}_{uuid.uuid4().hex[:12]}"
        self._paramMap: Dict[Param, Any] = {}
        self._defaultParamMap: Dict[Param, Any] = {}
        self._params_by_name: Dict[str, Param] = {}

    # -- declaration -----------------------------------------------------------------------------
    def _declare(self, name: str, doc: str, conv: Optional[Callable] = None) -> Param:
        p = Param(self, name, doc, conv)
        self._params_by_name[name] = p
        setattr(self, name, p)
        return p

    @property
    def params(self) -> List[Param]:
        return sorted(self._params_by_name.values(), key=lambda p: p.name)

    def _resetUid(self, uid: str) -> "Params":
        old = self.uid
        self.uid = uid
        for p in self._params_by_name.values():
            if p.parent == old:
                p.parent = uid
        self._paramMap = {self._params_by_name[p.name]: v for p, v in self._paramMap.items()}
        self._defaultParamMap = {self._params_by_name[p.name]: v for p, v in self._defaultParamMap.items()}
        return self

    # -- queries ---------------------------------------------------------------------------------
    def hasParam(self, name: str) -> bool:
        return isinstance(name, str) and name in self._params_by_name

    def getParam(self, name: str) -> Param:
        if name not in self._params_by_name:
            raise ValueError(f"Cannot find param with name {name}.")
        return self._params_by_name[name]

    def _resolve(self, param: Union[str, Param]) -> Param:
        if isinstance(param, Param):
            if param.name not in self._params_by_name:
                raise ValueError(f"Param {param} does not belong to {self.uid}.")
            return self._params_by_name[param.name]
        return self.getParam(param)

    def isSet(self, param: Union[str, Param]) -> bool:
        return self._resolve(param) in self._paramMap

    def hasDefault(self, param: Union[str, Param]) -> bool:
        return self.hasParam(param if isinstance(param, str) else param.name) and self._resolve(param) in self._defaultParamMap

    def isDefined(self, param: Union[str, Param]) -> bool:
        return self.isSet(param) or self.hasDefault(param)

    def getOrDefault(self, param: Union[str, Param]) -> Any:
        p = self._resolve(param)
        if p in self._paramMap:
            return self._paramMap[p]
        if p in self._defaultParamMap:
            return self._defaultParamMap[p]
        raise KeyError(f"Failed to find a default value for {p.name}")

    def extractParamMap(self, extra: Optional[Dict[Param, Any]] = None) -> Dict[Param, Any]:
        m = dict(self._defaultParamMap)
        m.update(self._paramMap)
        if extra:
            m.update(extra)
        return m

    # -- mutation --------------------------------------------------------------------------------
    def _set(self, **kwargs: Any) -> "Params":
        for name, value in kwargs.items():
            p = self.getParam(name)
            if value is not None:
                try:
                    value = p.typeConverter(value)
                except TypeError as e:
                    raise TypeError(f'Invalid param value given for param "{name}". {e}')
            self._paramMap[p] = value
        return self

    def _setDefault(self, **kwargs: Any) -> "Params":
        for name, value in kwargs.items():
            p = self.getParam(name)
            if value is not None:
                value = p.typeConverter(value)
            self._defaultParamMap[p] = value
        return self

    def set(self, param: Param, value: Any) -> None:
        self._set(**{param.name: value})

    def clear(self, param: Param) -> None:
        self._paramMap.pop(self._resolve(param), None)

    def copy(self, extra: Optional[Dict[Param, Any]] = None) -> "Params":
        that = _copy.copy(self)
        that._paramMap = dict(self._paramMap)
        that._defaultParamMap = dict(self._defaultParamMap)
        if extra:
            for p, v in extra.items():
                if not isinstance(p, Param):
                    raise TypeError(f"Expecting a valid instance of Param, but received: {p}")
                that._paramMap[that._resolve(p)] = v
        return that

    def _copyValues(self, to: "Params", extra: Optional[Dict[Param, Any]] = None) -> "Params":
        for p, v in self.extractParamMap(extra).items():
            if to.hasParam(p.name):
                if p in self._defaultParamMap and p not in self._paramMap and not (extra and p in extra):
                    to._defaultParamMap[to.getParam(p.name)] = v
                else:
                    to._paramMap[to.getParam(p.name)] = v
        return to


# ------------------------------------------------------------------------------------------------
class _CumlClass:
    """Mapping hooks every estimator / model implements (reference params.py:131-212)."""

    @classmethod
    def _param_mapping(cls) -> Dict[str, Optional[str]]:
        return {}

    @classmethod
    def _param_value_mapping(cls) -> Dict[str, Callable[[Any], Union[None, str, float, int]]]:
        return {}

    def _get_cuml_params_default(self) -> Dict[str, Any]:
        raise NotImplementedError()


class _CumlParams(_CumlClass, Params):
    """Keeps the Spark Params and the backend dict in sync (reference params.py:215-555)."""

    def __init__(self) -> None:
        super().__init__()
        self._cuml_params: Dict[str, Any] = {}
        self._num_workers: Optional[int] = None
        self._float32_inputs: bool = True

    @property
    def cuml_params(self) -> Dict[str, Any]:
        return self._cuml_params

    # -- workers ---------------------------------------------------------------------------------
    @property
    def num_workers(self) -> int:
        """One worker = one task = one GPU.  Inferred from the visible devices unless set
        (reference params.py:232-262: a request above the available GPUs raises in local mode)."""
        inferred = self._infer_num_workers()
        if self._num_workers is not None:
            if inferred < self._num_workers:
                raise ValueError(
                    f"The num_workers ({self._num_workers}) should be less than or equal to total GPUs ({inferred})"
                )
            return self._num_workers
        return inferred

    @num_workers.setter
    def num_workers(self, value: int) -> None:
        self._num_workers = value

    def _infer_num_workers(self) -> int:
        try:
            import torch

            n = torch.cuda.device_count()
            return max(n, 1)
        except Exception as e:  # same policy as the reference: never fail here
            print(e)
            return 1

    # -- mapping helpers ---------------------------------------------------------------------------
    def _initialize_cuml_params(self) -> None:
        self._cuml_params = self._get_cuml_params_default()
        for spark_param in self._param_mapping().keys():
            if self.hasParam(spark_param) and self.hasDefault(spark_param):
                self._set_cuml_param(spark_param, self.getOrDefault(spark_param))

    def _set_params(self, **kwargs: Any) -> "_CumlParams":
        param_map = self._param_mapping()
        for spark_param, cuml_param in param_map.items():
            if spark_param != cuml_param and spark_param in kwargs and cuml_param in kwargs:
                raise ValueError(f"'{cuml_param}' is an alias of '{spark_param}', set one or the other.")
        for k, v in kwargs.items():
            if k == "inputCol":
                self._set(**{"inputCol" if isinstance(v, str) else "inputCols": v})
            elif k == "featuresCol":
                self._set(**{"featuresCol" if isinstance(v, str) else "featuresCols": v})
            elif self.hasParam(k):
                self._set(**{str(k): v})
                self._set_cuml_param(k, v, silent=False)
            elif k in self._cuml_params:
                self._cuml_params[k] = v
                for spark_param, cuml_param in param_map.items():
                    if k == cuml_param and self.hasParam(spark_param):
                        try:
                            self._set(**{str(spark_param): v})
                        except TypeError:
                            pass  # the Spark-side converter may reject a backend-only value
            elif k == "num_workers":
                self._num_workers = v
            elif k == "float32_inputs":
                self._float32_inputs = v
            else:
                raise ValueError(f"Unsupported param '{k}'.")
        return self

    def clear(self, param: Param) -> None:
        super().clear(param)
        cuml_param = self._param_mapping().get(param.name)
        if cuml_param:
            self._cuml_params[cuml_param] = self._get_cuml_mapping_value(cuml_param, self.getOrDefault(param.name))

    def copy(self, extra: Optional[Dict[Param, Any]] = None) -> "_CumlParams":
        instance = super().copy(extra)
        cuml_params = dict(instance._cuml_params)
        if isinstance(extra, dict):
            for param, value in extra.items():
                name = instance._get_cuml_param(param.name, silent=False)
                if name is not None:
                    cuml_params[name] = instance._get_cuml_mapping_value(name, value)
        instance._cuml_params = cuml_params
        return instance

    def _copy_cuml_params(self, to: "_CumlParams") -> "_CumlParams":
        for k, v in self._cuml_params.items():
            if k in to._cuml_params:
                to._cuml_params[k] = v
        return to

    def _get_input_columns(self) -> Tuple[Optional[str], Optional[List[str]]]:
        """Order matters when several are defined: featuresCols beats featuresCol (reference :395-428)."""
        if self.hasParam("featuresCols") and self.isDefined("featuresCols"):
            return None, self.getOrDefault("featuresCols")
        if self.hasParam("featuresCol") and self.isDefined("featuresCol"):
            return self.getOrDefault("featuresCol"), None
        raise ValueError("Please set inputCol(s) or featuresCol(s)")

    def _get_cuml_param(self, spark_param: str, silent: bool = True) -> Optional[str]:
        param_map = self._param_mapping()
        if spark_param not in param_map:
            return None
        cuml_param = param_map[spark_param]
        if cuml_param is None:
            if not silent:
                raise ValueError(f"Spark Param '{spark_param}' is not supported by cuML.")
        elif cuml_param == "":
            if not silent:
                print(f"WARNING: Spark Param '{spark_param}' is not used by cuML.")
            cuml_param = None
        return cuml_param

    def _set_cuml_param(self, spark_param: str, spark_value: Any, silent: bool = True) -> None:
        cuml_param = self._get_cuml_param(spark_param, silent)
        if cuml_param is not None:
            try:
                self._set_cuml_value(cuml_param, spark_value)
            except ValueError:
                ref = cuml_param + " or " + spark_param if cuml_param != spark_param else spark_param
                raise ValueError(f"{ref} given invalid value {spark_value}")

    def _get_cuml_mapping_value(self, k: str, v: Any) -> Any:
        value_map = self._param_value_mapping()
        if k not in value_map:
            return v
        mapped = value_map[k](v)
        if mapped is None:
            raise ValueError(f"Value '{v}' for '{k}' param is unsupported")
        return mapped

    def _set_cuml_value(self, k: str, v: Any) -> None:
        self._cuml_params[k] = self._get_cuml_mapping_value(k, v)
