"""Name resolution by regular expression, with the semantics of LAB/utils/string.py:
`resolve_matching_names` (full-match regexes against a name list, optional key-order preservation, error on
unmatched or multiply-matched keys) and `resolve_matching_names_values` (dict regex -> value)."""
from __future__ import annotations

import re
from collections.abc import Sequence


def resolve_matching_names(keys, list_of_strings: Sequence[str], preserve_order: bool = False):
    if isinstance(keys, str):
        keys = [keys]
    index_list, names_list, key_idx_list = [], [], []
    matched_by = [None] * len(list_of_strings)
    keys_match_found = [[] for _ in keys]
    for target_idx, potential in enumerate(list_of_strings):
        for key_idx, regex in enumerate(keys):
            if re.fullmatch(regex, potential):
                if matched_by[target_idx] is not None:
                    raise ValueError(f"Multiple matches for '{potential}': '{matched_by[target_idx]}' and '{regex}'!")
                matched_by[target_idx] = regex
                index_list.append(target_idx)
                names_list.append(potential)
                key_idx_list.append(key_idx)
                keys_match_found[key_idx].append(potential)
    if preserve_order:
        order = sorted(range(len(index_list)), key=lambda i: (key_idx_list[i], index_list[i]))
        index_list = [index_list[i] for i in order]
        names_list = [names_list[i] for i in order]
    if not all(keys_match_found):
        msg = "\n".join(f"\t{k}: {v}" for k, v in zip(keys, keys_match_found))
        raise ValueError(f"Not all regular expressions are matched! Please check that the regular expressions are "
                         f"correct: \n{msg}\nAvailable strings: {list(list_of_strings)}\n")
    return index_list, names_list


def resolve_matching_names_values(data: dict, list_of_strings: Sequence[str], preserve_order: bool = False):
    if not isinstance(data, dict):
        raise TypeError(f"Input argument `data` should be a dictionary. Received: {data}")
    index_list, names_list, values_list, key_idx_list = [], [], [], []
    found = [[] for _ in data]
    for target_idx, potential in enumerate(list_of_strings):
        for key_idx, (regex, value) in enumerate(data.items()):
            if re.fullmatch(regex, potential):
                if target_idx in index_list:
                    raise ValueError(f"Multiple matches for '{potential}'!")
                index_list.append(target_idx)
                names_list.append(potential)
                values_list.append(value)
                key_idx_list.append(key_idx)
                found[key_idx].append(potential)
    if preserve_order:
        order = sorted(range(len(index_list)), key=lambda i: (key_idx_list[i], index_list[i]))
        index_list = [index_list[i] for i in order]
        names_list = [names_list[i] for i in order]
        values_list = [values_list[i] for i in order]
    if not all(found):
        raise ValueError(f"Not all regular expressions are matched! {dict(zip(data.keys(), found))}; "
                         f"available strings: {list(list_of_strings)}")
    return index_list, names_list, values_list
