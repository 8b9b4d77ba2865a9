"""Host-side utilities: recorders, logging, checkpoint IO, headless metrics, device timers."""
