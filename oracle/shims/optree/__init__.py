"""Oracle-only stand-in for `optree` (0.9.1 is pinned by the reference's
requirements.txt:82 but is not installable here: no network).

TEST INFRASTRUCTURE.  Only the oracle process puts this directory on sys.path so
that /root/reference/autoencoders/ensemble.py imports unmodified.  The reference
uses optree for pytree plumbing only (ensemble.py:51,56,60,65,108-109,166-173,
185-186) - no arithmetic - so a thin layer over torch.utils._pytree is exact.
optree sorts dict keys on flatten; torch keeps insertion order.  Both are
self-consistent between flatten and unflatten, which is all the reference needs.
"""
import torch.utils._pytree as _pt


def tree_flatten(tree):
    leaves, spec = _pt.tree_flatten(tree)
    return leaves, spec


def tree_unflatten(treespec, leaves):
    return _pt.tree_unflatten(list(leaves), treespec)


def tree_map(fn, tree, *rest):
    return _pt.tree_map(fn, tree, *rest)


def tree_map_(fn, tree, *rest):
    _pt.tree_map(fn, tree, *rest)
    return tree


def tree_leaves(tree):
    return _pt.tree_leaves(tree)
