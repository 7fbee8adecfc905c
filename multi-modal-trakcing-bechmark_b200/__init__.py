"""B200-native ViPT tracking forward."""
