// (filled in next)
