"""
Dynamic tasks on top of the pipeline scheduler (src/theia/task.py of the reference).
"""

from __future__ import annotations

import warnings
from typing import Callable

import numpy as np
from numpy.typing import NDArray

from theia_b200.pipeline import DynamicTask
from theia_b200.response import HistogramHitResponse

__all__ = ["ConvergeHistogramTask"]


def __dir__():
    return __all__


class ConvergeHistogramTask(DynamicTask):
    """
    Issues batches of a histogram until their mean has converged: the standard error of the
    mean of the per-batch bin sum must drop below ``atol + rtol * mean`` (src/theia/task.py:22-196).

    Parameters follow the reference: ``response, params, pipeline=None, *, initialBatchCount=4,
    extraBatchCount=2, maxBatchCount=50, atol=0.1, rtol=5e-5, finishedCallback=None``.
    """

    def __init__(self, response: HistogramHitResponse, params: dict, pipeline: str | None = None, *,
                 initialBatchCount: int = 4, extraBatchCount: int = 2, maxBatchCount: int = 50, atol: float = 0.1,
                 rtol: float = 5e-5, finishedCallback: Callable[["ConvergeHistogramTask"], None] | None = None) -> None:
        super().__init__(params, pipeline, initialBatchCount=initialBatchCount)
        if not response.retrieve:
            raise ValueError("Task requires response to retrieve results!")
        if initialBatchCount < 2:
            raise ValueError("initialBatchCount must be at least 2!")
        if extraBatchCount < 1:
            raise ValueError("extraBatchCount must be at least 1!")
        self._response = response
        self._extraBatchCount = extraBatchCount
        self._maxBatchCount = maxBatchCount
        self._atol, self._rtol = atol, rtol
        self._callback = finishedCallback
        self._nBatches = 0
        self._converged = False
        self._mean = np.zeros(response.nBins)  # running mean of the histograms
        self._sumMean = 0.0                    # running mean of the bin sums (Welford)
        self._sumM2 = 0.0                      # running sum of squared deviations of the bin sums

    @property
    def converged(self) -> bool:
        return self._converged

    @property
    def error(self) -> float:
        """Standard error of the mean of the bin sum, with the n - 1.5 small sample correction."""
        n = self._nBatches
        return float(np.sqrt(self._sumM2 / (n - 1.5)) / np.sqrt(n))

    @property
    def atol(self) -> float:
        return self._atol

    @property
    def rtol(self) -> float:
        return self._rtol

    @property
    def batchCountLimit(self) -> int:
        return self._maxBatchCount

    @property
    def response(self) -> HistogramHitResponse:
        return self._response

    @property
    def result(self) -> NDArray[np.float64]:
        """Mean histogram over all batches so far; check ``converged``."""
        return self._mean

    @property
    def totalBatches(self) -> int:
        return self._nBatches

    def onTaskFinished(self) -> None:
        if self._callback is not None:
            self._callback(self)

    def processBatch(self, config: int) -> int:
        hist = self._response.result(config)
        assert hist is not None
        self._nBatches += 1
        n = self._nBatches
        self._mean += (hist - self._mean) / n
        total = float(hist.sum())
        delta = total - self._sumMean
        self._sumMean += delta / n
        self._sumM2 += delta * (total - self._sumMean)
        if self.batchesRemaining > 0:
            return 0  # decide once everything issued so far has arrived
        if self.error <= self._atol + self._rtol * self._sumMean:
            self._converged = True
            return 0
        more = min(max(self._maxBatchCount - n, 0), self._extraBatchCount)
        if more == 0:
            warnings.warn(f"Failed to converge histogram (error: {self.error:.3e})!\n"
                          f"Pipeline: {self.pipeline}\nParams: {self.parameters}")
        return more
