"""TEST INFRASTRUCTURE ONLY -- a minimal stand-in for ``ete3.Tree`` so that the reference's own tree consumers
(``array2scaffolds.py:123-145 distance_matrix2tree``, ``:199-256 get_clusters``, ``:319-338 tree2scaffold``,
``bam2clusters.py:382-399 array2tree``) can be EXECUTED here by ``oracle/ref_exec.py``.

ete3 is a third-party dependency of the reference (PyPI ``ete3``, version unpinned in ``README.rst:17``), absent
from ``/root/reference`` and from this image.  Only the methods the reference calls are restated, following ete3
3.1's ``coretype/tree.py`` (TreeNode) semantics:

* ``Tree(newick=None, name=...)``; newick parsing of the ``(a:1.0,b:2.0):0.5;`` form (format 0, no quotes);
* ``add_child(child, dist=)``, ``remove_child``, ``get_children``, ``children``, ``up``, ``dist`` (default 1.0 for
  ``Tree()`` and for nodes read without a branch length, 0.0 for the root of a tree read from newick), ``name``
  ("" by default -- ete3's DEFAULT_NAME);
* ``is_leaf``, ``__len__`` (number of leaves), ``__iter__`` (leaves, pre-order, children left to right),
  ``traverse("postorder"|"preorder"|"levelorder")``, ``iter_leaves``, ``get_leaves``, ``get_leaf_names``;
* ``get_ancestors`` (parent first, root last);
* ``get_farthest_leaf()`` -> (leaf, distance): first leaf in pre-order with the strictly largest path length
  (``_get_farthest_and_closest_leaves``: ``if max_dist is None or total_d > max_dist``); a leaf returns (self, 0.0);
* ``write()`` -> newick format 0: ``name:dist`` for leaves, ``(...)support:dist`` for internal nodes (support 1),
  floats printed with ``%0.6g``, the root closes with ``;`` and carries no branch length.

The product never imports this file (it has its own iterative HCTree); it only serves to pin the oracle."""
from collections import deque

DEFAULT_NAME = ""       # ete3 3.x (ete2 used "NoName")
DEFAULT_DIST = 1.0      # TreeNode() without newick; a tree read from newick gets a root of dist 0.0


class Tree(object):
    def __init__(self, newick=None, name=None, dist=None, format=0):
        self.children = []
        self.up = None
        self.dist = DEFAULT_DIST if dist is None else float(dist)
        self.support = 1.0
        self.name = DEFAULT_NAME if name is None else name
        if newick is not None:
            self.dist = 0.0
            _read_newick(newick, self)

    # ---- structure -------------------------------------------------------------------------
    def add_child(self, child=None, name=None, dist=None, support=None):
        if child is None:
            child = Tree()
        if name is not None:
            child.name = name
        if dist is not None:
            child.dist = float(dist)
        if support is not None:
            child.support = support
        self.children.append(child)
        child.up = self
        return child

    def remove_child(self, child):
        self.children.remove(child)      # ValueError if absent, as ete3 raises
        child.up = None
        return child

    def get_children(self):
        return list(self.children)

    def is_leaf(self):
        return len(self.children) == 0

    def is_root(self):
        return self.up is None

    def get_ancestors(self):
        out, n = [], self.up
        while n is not None:
            out.append(n)
            n = n.up
        return out

    # ---- traversal -------------------------------------------------------------------------
    def _preorder(self):
        stack = [self]
        while stack:
            n = stack.pop()
            yield n
            stack.extend(reversed(n.children))

    def _postorder(self):
        stack = [(self, False)]
        while stack:
            n, seen = stack.pop()
            if seen or not n.children:
                yield n
            else:
                stack.append((n, True))
                for c in reversed(n.children):
                    stack.append((c, False))

    def _levelorder(self):
        q = deque([self])
        while q:
            n = q.popleft()
            yield n
            q.extend(n.children)

    def traverse(self, strategy="levelorder"):
        return {"preorder": self._preorder, "postorder": self._postorder, "levelorder": self._levelorder}[strategy]()

    def iter_leaves(self):
        for n in self._preorder():
            if not n.children:
                yield n

    def get_leaves(self):
        return list(self.iter_leaves())

    def get_leaf_names(self):
        return [n.name for n in self.iter_leaves()]

    def __iter__(self):
        return self.iter_leaves()

    def __len__(self):
        return sum(1 for _ in self.iter_leaves())

    def __bool__(self):
        return True

    __nonzero__ = __bool__

    # ---- distances ---------------------------------------------------------------------------
    def get_farthest_leaf(self, topology_only=False):
        """ete3's _get_farthest_and_closest_leaves: ONE running distance ``d`` that is increased on entering an
        internal node and decreased again on leaving it (iter_prepostorder), so the value seen at a leaf carries the
        rounding of that +=/-= history.  In an ultrametric tree all leaves are equally far up to rounding, so this
        arithmetic decides which leaf get_clusters prunes first: it is restated operation for operation."""
        if self.is_leaf():
            return self, 0.0
        best, best_d = None, None
        d = 0.0
        stack = [(False, c) for c in reversed(self.children)]
        while stack:
            post, n = stack.pop()
            if post:
                d -= 1.0 if topology_only else n.dist
            elif not n.children:
                total = d if topology_only else d + n.dist
                if best_d is None or total > best_d:
                    best, best_d = n, total
            else:
                d += 1.0 if topology_only else n.dist
                stack.append((True, n))
                stack.extend((False, c) for c in reversed(n.children))
        return best, best_d

    # ---- newick ------------------------------------------------------------------------------
    def write(self, format=0):
        out = []
        stack = [("node", self)]
        while stack:
            kind, n = stack.pop()
            if kind == "text":
                out.append(n)
                continue
            if n.children:
                out.append("(")
                tail = ")" if n is self else ")%0.6g:%0.6g" % (n.support, n.dist)
                stack.append(("text", tail))
                for k, c in enumerate(reversed(n.children)):
                    stack.append(("node", c))
                    if k != len(n.children) - 1:
                        stack.append(("text", ","))
            else:
                out.append("%s:%0.6g" % (n.name, n.dist))
        return "".join(out) + ";"


TreeNode = Tree


class TreeStyle(object):      # rendering is out of scope (no graphics stack): attributes are accepted and ignored
    pass


class _Faces(object):
    def TextFace(self, *a, **k):
        return None

    def add_face_to_node(self, *a, **k):
        return None


faces = _Faces()


def _read_newick(s, root):
    s = s.strip()
    if not s.endswith(";"):
        raise ValueError("newick must end with ';'")
    pos = 0
    cur = root
    root.name = ""
    n = len(s)
    first = True
    while pos < n:
        ch = s[pos]
        if ch == "(":
            if first:
                node = root
                first = False
                # root opens: its children follow
                child = Tree(name="")
                child.dist = 1.0
                node.children.append(child)
                child.up = node
                cur = child
            else:
                child = Tree(name="")
                child.dist = 1.0
                cur.children.append(child)
                child.up = cur
                cur = child
            pos += 1
        elif ch == ",":
            sib = Tree(name="")
            sib.dist = 1.0
            cur.up.children.append(sib)
            sib.up = cur.up
            cur = sib
            pos += 1
        elif ch == ")":
            cur = cur.up
            pos += 1
        elif ch == ";":
            break
        else:
            end = pos
            while end < n and s[end] not in "(),;":
                end += 1
            tok = s[pos:end]
            first = False
            name, _, dist = tok.partition(":")
            if cur.children:                 # internal node: the label is the support value (format 0)
                if name:
                    try:
                        cur.support = float(name)
                    except ValueError:
                        cur.name = name
            else:
                cur.name = name
            if dist:
                cur.dist = float(dist)
            pos = end
    return root
