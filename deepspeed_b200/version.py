"""Version of the B200-native framework (API level tracks reference ``version.txt:1`` = 0.19.6)."""
__version__ = "0.1.0"
__reference_api__ = "0.19.6"

version = __version__
git_hash = None
git_branch = None
