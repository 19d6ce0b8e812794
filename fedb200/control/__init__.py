"""Control plane: coordinator, worker servicer, primary/backup HA, CLI, fault injection."""
