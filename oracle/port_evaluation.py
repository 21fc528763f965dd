"""CPU restatement of the metric tables of evaluation.py:336-358 (inline closures of App.__init__,
not importable).  TEST INFRASTRUCTURE: only tests/ may import this.  Pure NumPy statistics over the
loss lists; pinned by hand-computable cases in tests/test_oracle_golden.py."""
import numpy as np


def compute_rte(t_losses):
    """:336-341."""
    rte_002 = np.sum(np.array(t_losses) <= 0.02) / len(t_losses)
    rte_005 = np.sum(np.array(t_losses) <= 0.05) / len(t_losses)
    rte_010 = np.sum(np.array(t_losses) <= 0.10) / len(t_losses)
    rte_015 = np.sum(np.array(t_losses) <= 0.15) / len(t_losses)
    return rte_002, rte_005, rte_010, rte_015


def compute_rre(r_losses):
    """:343-348."""
    r_losses = np.rad2deg(np.array(r_losses))
    return np.sum(r_losses <= 5) / len(r_losses), np.sum(r_losses <= 15) / len(r_losses), np.sum(r_losses <= 45) / len(r_losses)
