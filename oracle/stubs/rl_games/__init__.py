"""Stub of the two rl_games names the reference's RewardComponentObserver imports (rl/reward_observer.py:6-7).
TEST INFRASTRUCTURE ONLY: lets oracle/ref_harness.py run the UNMODIFIED reference observer in the build container
(rl_games is not installed there).  Restated from the published rl_games 1.6 sources."""
