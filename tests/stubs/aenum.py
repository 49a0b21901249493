"""Stand-in for the `aenum` dependency of vllm_omni/patch.py (absent from this image): only `extend_enum`,
which the reference uses to add members to vLLM enums at import time."""
import enum


def extend_enum(enumeration, name, value):
    if name in enumeration.__members__:
        return enumeration.__members__[name]
    base = enumeration._member_type_
    member = base.__new__(enumeration, value) if base is not object else object.__new__(enumeration)
    member._name_ = name
    member._value_ = value
    try:
        member._sort_order_ = len(enumeration._member_names_)
    except AttributeError:
        pass
    enumeration._member_names_.append(name)
    enumeration._member_map_[name] = member
    try:
        enumeration._value2member_map_[value] = member
    except TypeError:
        pass
    return member
