"""Short alias: `import b200feat` == `import expressive_speech_translation_b200` (SURVEY.md section 8b uses this name)."""
from expressive_speech_translation_b200 import *  # noqa: F401,F403
from expressive_speech_translation_b200 import __version__  # noqa: F401
