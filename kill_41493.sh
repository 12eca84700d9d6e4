ssh test@127.0.0.1 "kill -9 2791"
ssh test@127.0.0.1 "kill -9 2792"
ssh test@127.0.0.1 "kill -9 2793"
kill -9 2109
