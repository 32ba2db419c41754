"""Cluster-scheduling simulator: the reference's live tick-based "Horus" path (fifo / horus /
horus+ / gandiva over yarn / horus placement) and the event-driven Tiresias schedulers
(sjf, lpjf, shortest*, dlas*, gittins, ...) behind one CLI. See SURVEY.md §1 for the layer map."""
