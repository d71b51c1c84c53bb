"""Task layer of the drop-in (reference: plant3dvision/tasks/)."""
