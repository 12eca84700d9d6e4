ssh test@127.0.0.1 "kill -9 2490"
ssh test@127.0.0.1 "kill -9 2491"
kill -9 2109
