from dataclasses import dataclass
@dataclass
class GitStatus:
    commit: str = "unknown"
    has_unstaged_changes: bool = False
    has_staged_uncommitted_changes: bool = False
    has_untracked_files: bool = False
    @property
    def is_clean(self): return True
def git_status(*a, **k):
    return GitStatus()
