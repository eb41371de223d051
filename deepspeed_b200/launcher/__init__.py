"""Launchers: ``deepspeed`` CLI (runner), per-node process spawner (launch), multi-node command builders."""
