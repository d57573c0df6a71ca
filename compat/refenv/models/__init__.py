"""``models.sr_4k_2x`` / ``models.network_swinir`` of the reference scripts, served by this framework's model zoo."""
