ssh test@127.0.0.1 "kill -9 2438"
ssh test@127.0.0.1 "kill -9 2439"
ssh test@127.0.0.1 "kill -9 2440"
kill -9 2109
