"""Batched callers of the transform, shaped like the reference's own loops.

* ``classify_batch`` replaces the per-row loop of the CSV tab (``app_ui.py:144-150``: one Spark job pair per
  row) with ONE transform call for the whole frame.
* ``MicroBatcher`` replaces the per-message Kafka loop (``app_ui.py:195-226``: ``poll(1.0)`` -> classify ->
  ``produce``) with poll-up-to-``max_batch``-or-``max_wait`` -> one transform call -> one output record per message.
  It takes any consumer / producer with confluent_kafka's ``poll`` / ``produce`` / ``flush`` shape; the transport
  itself (librdkafka, brokers) stays out of scope.  The LLM explanation of the reference is not part of the record.
"""
from __future__ import annotations

import json
import time
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence

LABEL_MAPPING = {0: "Non-Fraudulent (Safe)", 1: "Potentially Fraudulent"}


def _scores(model, texts: Sequence[str]):
    out = model.transform(list(texts), output="host")
    names = model._plan.classifier
    pred = out[names.getOrDefault("predictionCol")]
    prob = out[names.getOrDefault("probabilityCol")]
    return pred, prob


def classify_batch(model, dialogues: Iterable[str]) -> List[Dict[str, Any]]:
    """-> one dict per dialogue with the columns the reference's batch tab builds (``app_ui.py:146-150``)."""
    texts = ["" if t is None else str(t) for t in dialogues]
    pred, prob = _scores(model, texts)
    return [{"dialogue": t, "predicted_label": LABEL_MAPPING.get(int(p), p), "prediction": float(p), "confidence": float(q[1])}
            for t, p, q in zip(texts, pred, prob)]


class MicroBatcher:
    """Accumulate Kafka-shaped messages and classify them one batch at a time.

    ``consumer.poll(timeout)`` must return ``None`` or a message with ``value()`` (JSON bytes ``{"text": ...}``,
    as the reference's producers send), ``key()`` and ``error()``; ``producer.produce(topic=, key=, value=)`` and
    ``producer.flush()`` are called once per batch."""

    def __init__(self, model, consumer, producer=None, output_topic: Optional[str] = None, max_batch: int = 4096,
                 max_wait_s: float = 0.05, poll_timeout_s: float = 0.01, on_error: Optional[Callable[[Any], None]] = None):
        self.model, self.consumer, self.producer, self.topic = model, consumer, producer, output_topic
        self.max_batch, self.max_wait_s, self.poll_timeout_s, self.on_error = int(max_batch), float(max_wait_s), float(poll_timeout_s), on_error
        self.batches = 0
        self.messages = 0

    def _collect(self):
        keys, texts = [], []
        deadline = None
        while len(texts) < self.max_batch:
            msg = self.consumer.poll(self.poll_timeout_s)
            now = time.monotonic()
            if msg is None:
                if deadline is None or now >= deadline:
                    break
                continue
            if msg.error():
                if self.on_error is not None:
                    self.on_error(msg.error())
                    continue
                raise RuntimeError(f"consumer error: {msg.error()}")
            try:
                payload = json.loads(msg.value().decode("utf-8"))
                text = payload["text"]
            except (ValueError, KeyError, AttributeError) as e:
                if self.on_error is not None:
                    self.on_error(e)
                    continue
                raise
            keys.append(msg.key())
            texts.append(text)
            if deadline is None:
                deadline = now + self.max_wait_s
            elif now >= deadline:
                break
        return keys, texts

    def step(self) -> List[Dict[str, Any]]:
        """One micro-batch: poll, classify, produce.  Returns the output records (empty when nothing arrived)."""
        keys, texts = self._collect()
        if not texts:
            return []
        pred, prob = _scores(self.model, texts)
        records = []
        for k, t, p, q in zip(keys, texts, pred, prob):
            rec = {"prediction": float(p), "confidence": float(q[1]), "original_text": t}
            records.append({"id": k, **rec})
            if self.producer is not None:
                self.producer.produce(topic=self.topic, key=k, value=json.dumps(rec))
        if self.producer is not None:
            self.producer.flush()
        self.batches += 1
        self.messages += len(texts)
        return records

    def run(self, should_stop: Callable[[], bool], idle_sleep_s: float = 0.0) -> None:
        while not should_stop():
            if not self.step() and idle_sleep_s:
                time.sleep(idle_sleep_s)
