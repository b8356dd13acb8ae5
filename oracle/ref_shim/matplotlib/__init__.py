"""matplotlib stand-in (render-only dependency of the reference; never exercised)."""
