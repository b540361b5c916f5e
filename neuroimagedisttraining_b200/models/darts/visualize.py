"""Genotype -> graph description.  The reference renders cells with the ``graphviz`` package (``darts/visualize.py:1-57``); here the
DOT source is produced directly (any Graphviz install renders it) plus a plain-text listing, so no optional dependency is needed."""
from .genotypes import Genotype


def cell_edges(cell):
    """[(source node, target node, op)] with nodes named ``c_{k-2}``, ``c_{k-1}``, ``0`` .. ``steps-1``."""
    assert len(cell) % 2 == 0
    edges = []
    for i in range(len(cell) // 2):
        for op, j in cell[2 * i:2 * i + 2]:
            src = "c_{k-2}" if j == 0 else "c_{k-1}" if j == 1 else str(j - 2)
            edges.append((src, str(i), op))
    return edges


def to_dot(cell, name="cell") -> str:
    steps = len(cell) // 2
    lines = [f"digraph {name} {{", "  rankdir=LR;", '  node [style=filled, shape=rect, fontname="helvetica"];',
             '  "c_{k-2}" [fillcolor=darkseagreen2]; "c_{k-1}" [fillcolor=darkseagreen2];']
    lines += [f'  "{i}" [fillcolor=lightblue];' for i in range(steps)]
    lines += [f'  "{s}" -> "{t}" [label="{op}"];' for s, t, op in cell_edges(cell)]
    lines += ['  "c_{k}" [fillcolor=palegoldenrod];'] + [f'  "{i}" -> "c_{{k}}";' for i in range(steps)] + ["}"]
    return "\n".join(lines)


def describe(genotype: Genotype) -> str:
    out = []
    for kind in ("normal", "reduce"):
        out.append(f"{kind} cell (concat {list(getattr(genotype, kind + '_concat'))}):")
        out += [f"  {s:>8} --{op}--> {t}" for s, t, op in cell_edges(getattr(genotype, kind))]
    return "\n".join(out)


def plot(cell, filename: str) -> str:
    """Writes ``<filename>.dot`` and returns its path."""
    path = filename if filename.endswith(".dot") else filename + ".dot"
    with open(path, "w") as f:
        f.write(to_dot(cell, name="cell"))
    return path
